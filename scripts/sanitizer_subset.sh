#!/bin/bash
# compute-sanitizer over the reduced-size part of the GPU suite (the 240 s / T = 5760 cases would run for hours under
# the tool): one tool per gpurun call, as the profiling recipe asks.  usage: scripts/sanitizer_subset.sh memcheck|racecheck
TOOL=${1:-memcheck}
SEL='not full_size and not fullsize and not 240s and not full_length and not parseval and not tapered and not graph and not philox and not deferred_sampler and not config2 and not long_clip'
timeout 1500 compute-sanitizer --tool $TOOL --error-exitcode 99 --print-limit 20 \
  python -m pytest tests/test_gpu_frontend.py tests/test_gpu_chroma.py tests/test_gpu_post.py tests/test_gpu_hpss_filters.py \
  tests/test_gpu_a2f.py tests/test_gpu_traj.py tests/test_gpu_builders.py tests/test_gpu_maua.py tests/test_known_answers.py \
  -m gpu -q -x -k "$SEL" 2>&1 | tail -40
echo "exit code: ${PIPESTATUS[0]}"
