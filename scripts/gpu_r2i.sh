#!/bin/bash
# Round 2, call I: why is the (cache write, attention) loop of bench.py slower than the decode kernel alone?
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for t in 0 512; do
  python bench.py --steps 300 --warmup 5 --no-prefill --no-cpu-baseline --no-comparators --tune $t > gpurun_out/r2i_bench_t$t.log 2>&1
  python - <<PY
import json
d = json.loads([l for l in open("gpurun_out/r2i_bench_t$t.log") if l.startswith("{")][-1])
print("tune $t", d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], "roofline", d["roofline"]["avg_launch_ms"], d["extras"]["other_decode_configs"])
PY
done
python - <<'PY'
import sys, torch, time
sys.path.insert(0, ".")
import bench
from vllm_triton_backend_b200 import _cabi
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
s = torch.cuda.current_stream()
def t(fn, n=300):
    for _ in range(10): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); 
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
for wl, hq, hkv in (("c2", 32, 8), ("c4", 8, 1)):
    for tune in (0, 512):
        _cabi.set_tuning(0, 0, 0, tune)
        hp = bench.HotPath(bench.WORKLOADS[wl], hq, hkv, dev, 0)
        print(wl, hkv, "tune", tune, "attention only %.1f us" % t(hp.attention), "cache write only %.1f us" % t(hp.cache_write), "pair %.1f us" % t(hp.step))
        t0 = time.perf_counter()
        for _ in range(2000): hp.step()
        host = (time.perf_counter() - t0) / 2000 * 1e6
        torch.cuda.synchronize()
        print("   host time per step (launch calls only) %.1f us" % host)
        hp.free(); del hp
PY
