# Attribution of the training chains' cost over inference (446 vs 296 us per fine forward launch): rebuild the library on
# the GPU box with one piece of the saved-image path removed or replaced at a time (wrong gradients, right timing).
#   -DCNRF_EXP_NO_STG            the store warp reads the tile out of shared memory but does not store it
#   -DCNRF_EXP_NO_COPY           no copy at all (the hand-shake stays)
#   -DCNRF_EXP_COALESCED_DIRECT  no copy; the epilogue stores its packed 32-column blocks itself with a warp-coalesced
#                                address pattern ([16-byte chunk][row]: what a chunk-major image layout would allow)
mkdir -p gpurun_out
cp -r code-nerf_b200/lib /tmp/lib_keep
variants=("" "-DCNRF_EXP_NO_STG" "-DCNRF_EXP_NO_COPY" "-DCNRF_EXP_COALESCED_DIRECT")
for v in "${variants[@]}"; do
  CNRF_NVCC_EXTRA="$v" python code-nerf_b200/build.py --force > /dev/null 2>&1 || { echo "build failed: $v"; continue; }
  echo "variant [$v]"
  timeout 120 python scripts/time_chain_kernels.py 20 2>&1 | grep "fine-pass"
done > gpurun_out/fwd_store_attribution.txt 2>&1
rm -rf code-nerf_b200/lib; mv /tmp/lib_keep code-nerf_b200/lib
cat gpurun_out/fwd_store_attribution.txt
