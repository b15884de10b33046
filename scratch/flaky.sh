mkdir -p gpurun_out
for i in 1 2 3 4; do
  for pf in 0 512; do
    MB200_BODY_PREFETCH=$pf timeout 120 python -m pytest "tests/test_gpu_collision.py::test_bbpgd_converged" -q -s 2>&1 | grep -E "iters|assert|Error|passed|failed|^E " | head -12 | sed "s/^/[pf=$pf run $i] /"
  done
done
nproc
