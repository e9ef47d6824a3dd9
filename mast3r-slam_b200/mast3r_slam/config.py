"""Global configuration dict, as the reference's mast3r_slam/config.py:1-54: a module-level
`config` that callers index (`config["matching"]["max_iter"]`), filled by `load_config(path)`
from YAML files that may name a parent with `inherit:`.  Defaults below are the values of the
reference's config/base.yaml for the sections the hot path reads (:8-50)."""
import copy
import re

config = {
    "use_calib": False,
    "matching": {"max_iter": 10, "lambda_init": 1e-8, "convergence_thresh": 1e-6,
                 "dist_thresh": 1e-1, "radius": 3, "dilation_max": 5},
    "tracking": {"min_match_frac": 0.05, "max_iters": 50, "C_conf": 0.0, "Q_conf": 1.5,
                 "rel_error": 1e-3, "delta_norm": 1e-3, "huber": 1.345, "match_frac_thresh": 0.333,
                 "sigma_ray": 0.003, "sigma_dist": 10.0, "sigma_pixel": 1.0, "sigma_depth": 10.0,
                 "sigma_point": 0.05, "pixel_border": -10, "depth_eps": 1e-6,
                 "filtering_mode": "weighted_pointmap", "filtering_score": "median"},
    "local_opt": {"pin": 1, "window_size": 1e6, "C_conf": 0.0, "Q_conf": 1.5, "min_match_frac": 0.1,
                  "pixel_border": -10, "depth_eps": 1e-6, "max_iters": 10, "sigma_ray": 0.003,
                  "sigma_dist": 10.0, "sigma_pixel": 1.0, "sigma_depth": 10.0, "sigma_point": 0.05,
                  "delta_norm": 1e-8, "use_cuda": True},
}
_DEFAULTS = copy.deepcopy(config)

# YAML 1.1 reads `1e-8` as a string; the reference installs a resolver so it is a float
_FLOAT = re.compile(r"^[-+]?(\d[\d_]*\.?[\d_]*|\.[\d_]+)([eE][-+]?\d+)?$|^[-+]?\.(inf|Inf|INF)$|^\.(nan|NaN|NAN)$")


def _deep_update(dst, src):
    for k, v in src.items():
        if isinstance(v, dict):
            node = dst.setdefault(k, {})
            if not isinstance(node, dict):
                node = dst[k] = {}
            _deep_update(node, v)
        else:
            dst[k] = v
    return dst


def _read(path):
    import yaml

    class Loader(yaml.SafeLoader):
        pass

    Loader.add_implicit_resolver("tag:yaml.org,2002:float", _FLOAT, list("-+0123456789."))
    with open(path, "r") as f:
        node = yaml.load(f, Loader=Loader) or {}
    parent = node.get("inherit")
    return _deep_update(_read(parent) if parent else {}, node)


def load_config(path="config/base.yaml"):
    """Read `path` (and its `inherit:` chain, parents first) into the global dict."""
    return _deep_update(config, _read(path))


def reset_config():
    config.clear()
    config.update(copy.deepcopy(_DEFAULTS))
    return config
