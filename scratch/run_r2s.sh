#!/bin/bash
# one compute-sanitizer tool (memcheck) over the small tests of this round's new code paths
mkdir -p gpurun_out
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 66 --print-limit 20 \
  python -m pytest -m gpu -x -q tests/test_zz_no_alloc.py tests/test_gpu_mg.py tests/test_gpu_closed_form.py \
  "tests/test_gpu_raster.py::test_batched_views_equal_single_passes" "tests/test_gpu_raster.py::test_chunk_queue_overflow_falls_back" \
  "tests/test_gpu_tier_b.py::test_shade_ldr_equals_shade_then_tonemap" "tests/test_gpu_tier_b.py::test_shade_replicated_stores_equal_the_plain_frame" \
  "tests/test_gpu_tier_b.py::test_shade_host_light_list_kernels" "tests/test_gpu_tier_b.py::test_prepared_source_equals_the_plain_entry_points" \
  "tests/test_gpu_tier_a.py::test_batched_fused_views_equal_two_passes" > gpurun_out/memcheck.txt 2>&1
echo "memcheck exit $?"
grep -E "ERROR SUMMARY|Invalid|passed|failed|error" gpurun_out/memcheck.txt | tail -15
