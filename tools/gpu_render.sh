timeout 900 python -m pytest tests/test_render.py tests/test_variations.py -m gpu -x -q 2>&1 | tail -12
timeout 600 python bench.py --workload hammer-v0:8192 --pixels --steps 60 --warmup 10 --no-cpu-baseline --also '' > gpurun_out/r2_bench_hammer_pixels.json 2>gpurun_out/r2_bench_hammer_pixels.err; tail -c 600 gpurun_out/r2_bench_hammer_pixels.err; cut -c1-400 gpurun_out/r2_bench_hammer_pixels.json
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,dram__bytes_write.sum,smsp__inst_executed.sum -k regex:render_kernel -s 20 -c 5 --csv --log-file gpurun_out/r2_render_kernel.csv python bench.py --workload hammer-v0:8192 --pixels --steps 30 --warmup 3 --no-cpu-baseline --also '' > /dev/null 2>&1
python - <<'PY'
import torch, numpy as np, sys
sys.path.insert(0, '.')
from mj_envs_b200.vec_env import AdroitVecEnv
env = AdroitVecEnv("hammer-v0", 4, seed=0); env.reset()
img = env.render().cpu().numpy()
np.save("gpurun_out/r2_render_sample_hammer.npy", img)
PY
