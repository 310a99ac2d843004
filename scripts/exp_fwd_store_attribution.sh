# Attribution of the training forward's cost over the inference forward (466 vs 296 us per fine launch): rebuild the
# library on the GPU box with one piece of the saved-image path removed at a time (wrong gradients, right timing).
mkdir -p gpurun_out
cp -r code-nerf_b200/lib /tmp/lib_keep
for v in "" "-DCNRF_EXP_NO_STG" "-DCNRF_EXP_NO_COPY"; do
  CNRF_NVCC_EXTRA="$v" python code-nerf_b200/build.py --force > /dev/null 2>&1 || { echo "build failed: $v"; continue; }
  echo "variant [$v]"
  timeout 120 python scripts/time_chain_kernels.py 20 2>&1 | grep "fine-pass"
done > gpurun_out/fwd_store_attribution.txt 2>&1
rm -rf code-nerf_b200/lib; mv /tmp/lib_keep code-nerf_b200/lib
cat gpurun_out/fwd_store_attribution.txt
