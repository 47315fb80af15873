# Round-end check on the GPU box: full -m gpu tests, then short bench lines into gpurun_out/.
mkdir -p gpurun_out
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for L in 1 2 4 8; do python bench.py --task finger-spin --lanes $L --steps 20 --warmup 5 --no-cpu 2>&1 | tail -1 > gpurun_out/r1o_finger_g$L.json; done
python bench.py --task cheetah-run --steps 50 --warmup 5 --no-cpu 2>&1 | tail -1 > gpurun_out/r1o_cheetah.json
python bench.py --task humanoid-run --steps 20 --warmup 5 --no-cpu 2>&1 | tail -1 > gpurun_out/r1o_humanoid.json
