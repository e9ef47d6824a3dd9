# Sweep of the macro-tile Cholesky's chunk geometry (compile-time constants): rebuilds the library per point and
# prints the N=512 solve step (CUDA events, chol_trace.py).  Run under gpurun; restores the sources afterwards.
set -e
cp mast3r-slam_b200/csrc/m3r_gn.cuh /tmp/gn.cuh.orig; cp mast3r-slam_b200/csrc/m3r_gn_chol64.cu /tmp/c64.cu.orig
for cfg in ${M3R_SWEEP:-"8 16 8" "12 16 8" "8 8 8" "8 32 8" "12 32 8" "8 16 4" "8 16 12" "12 16 12"}; do
  set -- $(echo $cfg | tr "," " ")
  cp /tmp/gn.cuh.orig mast3r-slam_b200/csrc/m3r_gn.cuh; cp /tmp/c64.cu.orig mast3r-slam_b200/csrc/m3r_gn_chol64.cu
  sed -i "s/constexpr int kWindow = 8;/constexpr int kWindow = $1;/" mast3r-slam_b200/csrc/m3r_gn_chol64.cu
  sed -i "s/constexpr int kGnKCOld = 16;/constexpr int kGnKCOld = $2;/; s/constexpr int kGnRecent = 4;/constexpr int kGnRecent = $3;/; s/constexpr int kGnKC = 4;/constexpr int kGnKC = ${4:-4};/; s/constexpr int kGnTailChain = 2;/constexpr int kGnTailChain = ${5:-2};/" mast3r-slam_b200/csrc/m3r_gn.cuh
  python mast3r-slam_b200/build.py > /dev/null 2>&1
  r=$(timeout 100 python profiles/chol_trace.py 512 /tmp/sweep_trace.txt 2>/dev/null | head -1)
  f=$(grep factoris /tmp/sweep_trace.txt)
  echo "window $1 old-chunk $2 recent $3 chunk ${4:-4} chain-tail ${5:-2} : $r ; $f"
done
cp /tmp/gn.cuh.orig mast3r-slam_b200/csrc/m3r_gn.cuh; cp /tmp/c64.cu.orig mast3r-slam_b200/csrc/m3r_gn_chol64.cu
python mast3r-slam_b200/build.py > /dev/null 2>&1
