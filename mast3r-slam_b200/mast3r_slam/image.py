"""Mirror of /root/reference/mast3r_slam/image.py for the hot path."""
import mast3r_slam_backends as _be


def img_gradient(img):
    """image.py:5-38: reflect-padded depthwise 3x3 gradients of img (b,c,h,w) -> (gx, gy).
    CUDA float32 contiguous input; the arithmetic is one kernel (m3r_img_gradient).  The matching
    pipeline does not call it — m3r_match_prep computes the same taps fused with the ray
    normalisation — it exists for callers of the reference function."""
    return _be.img_gradient(img)
