#!/usr/bin/env python
"""Benchmark of the paged-KV attention hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c4|c3]

N=1 workload = BASELINE.json configs[1] ("C2"): Llama-3.1-8B-shaped paged decode,
bf16, 32 q / 8 kv heads, D=128, batch 64, ctx 4096, block_size 16.  A step is one
pass of the hot path over that batch: KV-cache write of the 64 new tokens (K1)
followed by split-KV decode attention (K2 + segment reduce).  The metric is decode
KV GB/s = algorithmic bytes / device time.

N>1 (torchrun, one rank per GPU): the path shards by KV head with no collective
(vLLM tensor parallelism).  The workload is then BASELINE.json's
multi-GPU configuration, C4: Llama-3.1-70B-shaped decode, 64q/8kv heads, FP8-e4m3
KV, batch 64, ctx 32k, its 8 KV heads split over the ranks ("strong": total work
fixed), `value` = total algorithmic bytes / max time over ranks.  One C2-shaped
replica per rank ("weak", round 1's line) is reported under extras; `--workload`
overrides either default.

`--impl reference` times the CPU restatement of the reference path (oracle/, the
reference being Python+Triton has no CPU binary to run on the GPU box) on the
same config with all host threads; rank 0 only.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402

WORKLOADS = {
    # name: (batch, ctx, Hq, Hkv, D, block_size, kv dtype)
    "c2": dict(batch=64, ctx=4096, hq=32, hkv=8, d=128, bs=16, kv="bf16",
               desc="C2: Llama-3.1-8B-shaped paged decode bf16, 32q/8kv, D=128, batch 64, ctx 4096, block_size 16"),
    "c4": dict(batch=64, ctx=32768, hq=64, hkv=8, d=128, bs=16, kv="fp8",
               desc="C4: Llama-3.1-70B-shaped decode, 64q/8kv, FP8-e4m3 KV, batch 64, ctx 32768, KV heads sharded"),
}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm_gbs=float(d["hbm_gbs"]), bf16_tflops=float(d["bf16_tflops"]),
                    bf16_tflops_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
    return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, bf16_tflops_sustained=1400.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu_index = gpu_index
        self.lines = []
        self.proc = None
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50", "-i", str(self.gpu_index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.t0 = time.time()

        def pump():
            for line in self.proc.stdout:
                self.lines.append((time.time(), line.strip()))

        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()

    def mark(self):
        return time.time()

    def stop(self, t_begin=None, t_end=None):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        rows = []
        for ts, line in self.lines:
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                rows.append((ts, float(parts[1]), float(parts[2]), parts[4:]))
            except ValueError:
                continue
        if not rows:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["no samples"])
        inside = [r for r in rows if (t_begin is None or r[0] >= t_begin) and (t_end is None or r[0] <= t_end + 0.12)]
        use = inside or rows[-3:]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in use:
            flags = r[3]
            for name, val in zip(names, flags[1:5]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return dict(sm_mhz=statistics.median(r[1] for r in use), sm_max_mhz=max(r[2] for r in use),
                    reasons=sorted(reasons), samples=len(use), samples_in_region=len(inside))


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def algorithmic_bytes(w, hq, hkv, kv_bytes):
    from vllm_triton_backend_b200.roofline import cache_write_bytes, decode_bytes

    seq_lens = [w["ctx"]] * w["batch"]
    dec = decode_bytes(seq_lens, hq, hkv, w["d"], w["bs"], kv_bytes)
    cw = cache_write_bytes(w["batch"], hkv, w["d"], 2, kv_bytes)
    return dec, cw


def make_device_workload(w, hq, hkv, device, seed):
    """Synthetic C2/C4-shaped state resident in HBM: KV cache (random permutation
    block table so L2 cannot alias pages), one new token per sequence."""
    g = torch.Generator(device=device).manual_seed(seed)
    B, ctx, D, bs = w["batch"], w["ctx"], w["d"], w["bs"]
    pages_per_seq = ctx // bs
    NB = B * pages_per_seq + 1024
    fp8 = w["kv"] == "fp8"
    # (2, NB, bs, Hkv, D): the layout B200AttentionBackend.get_kv_cache_shape declares
    if fp8:
        kv = torch.empty(2, NB, bs, hkv, D, dtype=torch.uint8, device=device)
        chunk = 1 << 26
        flat = kv.view(-1)
        for o in range(0, flat.numel(), chunk):
            n = min(chunk, flat.numel() - o)
            x = (torch.rand(n, device=device, generator=g) * 2 - 1) / 0.5
            flat[o:o + n] = x.to(torch.float8_e4m3fn).view(torch.uint8)
    else:
        kv = torch.empty(2, NB, bs, hkv, D, dtype=torch.bfloat16, device=device)
        chunk = 1 << 26
        flat = kv.view(-1)
        for o in range(0, flat.numel(), chunk):
            n = min(chunk, flat.numel() - o)
            flat[o:o + n] = (torch.rand(n, device=device, generator=g) * 2 - 1).to(torch.bfloat16)
    perm = torch.randperm(NB, device=device, generator=g)[: B * pages_per_seq].to(torch.int32)
    block_table = perm.view(B, pages_per_seq).contiguous()
    seq_lens = torch.full((B,), ctx, dtype=torch.int32, device=device)
    cu = torch.arange(B + 1, dtype=torch.int32, device=device)
    # the new token of each sequence sits at position ctx-1
    last = block_table[:, -1].to(torch.int64) * bs + (ctx - 1) % bs
    q = (torch.rand(B, hq, D, device=device, generator=g) * 2 - 1).to(torch.bfloat16)
    k_new = (torch.rand(B, hkv, D, device=device, generator=g) * 2 - 1).to(torch.bfloat16)
    v_new = (torch.rand(B, hkv, D, device=device, generator=g) * 2 - 1).to(torch.bfloat16)
    return dict(kv=kv, block_table=block_table, seq_lens=seq_lens, cu=cu, slot_mapping=last, q=q, k_new=k_new, v_new=v_new)



def _mixed_bounds(m, peaks):
    """A mixed batch has a tensor-bound share (the prompts' FLOPs) and an HBM-bound share (the KV the
    single-token rows stream).  Lower bounds on its time from the two measured peaks: the larger of the two
    if they overlap perfectly, their sum when they run one after the other."""
    t_flop = m["flops"] / (peaks["bf16_tflops"] * 1e12) * 1e3
    t_byte = m["decode_row_kv_bytes"] / (peaks["hbm_gbs"] * 1e9) * 1e3
    return {"decode_rows": m["decode_rows"], "decode_row_kv_gb": round(m["decode_row_kv_bytes"] / 1e9, 4),
            "bound_ms_overlapped": round(max(t_flop, t_byte), 4), "bound_ms_serial": round(t_flop + t_byte, 4),
            "frac_of_overlapped_bound": round(max(t_flop, t_byte) / m["ms"], 4),
            "frac_of_serial_bound": round((t_flop + t_byte) / m["ms"], 4)}


def measure_prefill(device, world):
    """Tensor-bound half of the metric: prefill attention TFLOP/s on one full 8k causal
    prompt and on the C3 chunked-prefill mixed batch (Llama-3.1-8B shape), through
    ops.unified_attention.  CUDA events around >= 50 ms of launches after warm-up."""
    import types

    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import kbench

    out = {}
    for shape in ("full8k", "mixed"):
        a = types.SimpleNamespace(shape=shape, hq=0, hkv=0, bs=0, fp8=False, layout="planes", force=None)
        fn, flops, cfg = kbench.prefill_case(a, device)
        ms, n = kbench.time_fn(fn, min_ms=50.0)
        out[shape] = dict(ms=ms, tflops=flops / ms / 1e9, tokens=cfg["T"], iters=n, flops=flops,
                          decode_rows=cfg["decode_rows"], decode_row_kv_bytes=cfg["decode_row_kv_bytes"])
        del fn
        torch.cuda.empty_cache()
    return out


class _Layer:
    """The attributes of vLLM's Attention layer the impl reads."""

    def __init__(self, device, k_scale=1.0, v_scale=1.0):
        self._k_scale = torch.tensor(k_scale, dtype=torch.float32, device=device)
        self._v_scale = torch.tensor(v_scale, dtype=torch.float32, device=device)
        self._q_scale = torch.tensor(1.0, dtype=torch.float32, device=device)


class HotPath:
    """One workload resident in HBM and the calls a step makes through the plugin surface:
    `B200AttentionImpl.do_kv_cache_update` (K1) then `.forward` (K2), with the step's metadata - including the
    split-KV work plan the metadata builder computes once per step - in a `B200AttentionMetadata`."""

    def __init__(self, w, hq, hkv, device, seed):
        from vllm_triton_backend_b200 import ops
        from vllm_triton_backend_b200.backend.attn import B200AttentionImpl, B200AttentionMetadata

        self.w, self.hq, self.hkv, self.device = w, hq, hkv, device
        self.fp8 = w["kv"] == "fp8"
        self.st = st = make_device_workload(w, hq, hkv, device, seed)
        B, D = w["batch"], w["d"]
        self.layer = _Layer(device, 0.5 if self.fp8 else 1.0, 0.5 if self.fp8 else 1.0)
        self.impl = B200AttentionImpl(num_heads=hq, head_size=D, scale=D ** -0.5, num_kv_heads=hkv, alibi_slopes=None,
                                      sliding_window=None, kv_cache_dtype="fp8" if self.fp8 else "auto")
        self.workspace = torch.zeros(ops.workspace_bytes(B, B, hq, hkv, D), dtype=torch.uint8, device=device)
        kv_dtype = torch.float8_e4m3fn if self.fp8 else torch.bfloat16
        self.new_plan = lambda: ops.DecodePlan(device, B, hq, hkv, D, w["bs"], torch.bfloat16, kv_dtype)
        self.plan = self.new_plan()
        self.plan.build(st["cu"], st["seq_lens"], 1)  # what B200AttentionMetadataBuilder.build does every step
        self.Meta = B200AttentionMetadata
        self.meta = self.metadata(st["cu"], st["seq_lens"], st["block_table"], st["slot_mapping"], self.plan)
        self.out = torch.empty(B, hq, D, dtype=torch.bfloat16, device=device)
        kvb = 1 if self.fp8 else 2
        self.dec_bytes, self.cw_bytes = algorithmic_bytes(w, hq, hkv, kvb)
        self.step_bytes = self.dec_bytes + self.cw_bytes

    def metadata(self, cu, seq_lens, block_table, slots, plan):
        return self.Meta(num_actual_tokens=self.w["batch"], max_query_len=1, query_start_loc=cu, max_seq_len=self.w["ctx"],
                         seq_lens=seq_lens, block_table=block_table, slot_mapping=slots, workspace=self.workspace,
                         decode_plan=plan)

    def cache_write(self):
        st = self.st
        self.impl.do_kv_cache_update(self.layer, st["k_new"], st["v_new"], st["kv"], st["slot_mapping"])

    def attention(self):
        st = self.st
        self.impl.forward(self.layer, st["q"], st["k_new"], st["v_new"], st["kv"], self.meta, output=self.out)

    def step(self):
        self.cache_write()
        self.attention()

    def free(self):
        self.st.clear()
        self.meta = self.out = self.workspace = None
        torch.cuda.empty_cache()


def time_steps(hp, K, warmup, barrier, stream):
    """W untimed steps, then exactly K steps between two events on the launching stream; ms total."""
    for _ in range(warmup):
        hp.step()
    barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record(stream)
    for _ in range(K):
        hp.step()
    ev[1].record(stream)
    barrier()
    return ev[0].elapsed_time(ev[1])


def roofline_pass(hp, launches, stream):
    """Per-launch duration of the dominant kernel: `launches` attention launches, each bracketed by its own pair of
    events on the launching stream, in a pass of its own (an event record between two launches serialises them, so
    doing this inside the timed loop would change what `value` measures)."""
    for _ in range(5):
        hp.attention()
    torch.cuda.synchronize(hp.device)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(launches)]
    for a, b in evs:
        a.record(stream)
        hp.attention()
        b.record(stream)
    torch.cuda.synchronize(hp.device)
    return [a.elapsed_time(b) for a, b in evs]


def e2e_pass(hp, K, warmup, barrier, serial):
    """The same step through the plugin with HOST buffers: the step's inputs live in one pinned arena
    (vllm_triton_backend_b200.staging), go to the device with one copy, the split-KV plan is rebuilt from the
    uploaded metadata (as the builder does each step), and the output comes back with one copy."""
    from vllm_triton_backend_b200.staging import StepPipeline

    st, w = hp.st, hp.w
    B, D = w["batch"], w["d"]
    names = {"q": st["q"], "k": st["k_new"], "v": st["v_new"], "slots": st["slot_mapping"], "bt": st["block_table"],
             "sl": st["seq_lens"], "cu": st["cu"]}
    depth = 1 if serial else 2
    pipe = StepPipeline([(n, tuple(t.shape), t.dtype) for n, t in names.items()], (B, hp.hq, D), torch.bfloat16, hp.device,
                        depth=depth)
    plans = {}
    for slot in pipe.slots:
        for n, t in names.items():
            slot.staging.host[n].copy_(t.cpu())
        plans[id(slot.staging.dev)] = hp.new_plan()
    h2d = pipe.slots[0].staging.payload_bytes
    d2h = B * hp.hq * D * 2

    def compute(d, o):
        plan = plans[id(d)]
        plan.build(d["cu"], d["sl"], 1)
        m = hp.metadata(d["cu"], d["sl"], d["bt"], d["slots"], plan)
        hp.impl.do_kv_cache_update(hp.layer, d["k"], d["v"], st["kv"], d["slots"])
        hp.impl.forward(hp.layer, d["q"], d["k"], d["v"], st["kv"], m, output=o)

    checksum = 0.0

    def e2e_step():
        nonlocal checksum
        slot = pipe.acquire()  # blocks until the result of the step that last used this slot is on the host
        checksum += float(slot.host_out[0, 0, 0])  # the caller reads every step's result
        pipe.run(slot, compute)

    for _ in range(warmup):
        e2e_step()
    pipe.drain()
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        e2e_step()
    pipe.drain()
    barrier()
    return (time.perf_counter() - t0) * 1e3, h2d, d2h, depth


def same_gpu_comparators(hp, budget_ms=60.0):
    """The reference's own kernels on the same tensors of the same GPU (tests/triton_reference.py: the reference
    file from baseline/_ref when build() put it there, else the copy the installed vLLM ships)."""
    out = {}
    try:
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        import kbench
        import triton_reference as tr

        why = tr.available()
        if why:
            return {"triton_unified_attention": {"unavailable": why}}
        st, w = hp.st, hp.w
        B, D = w["batch"], w["d"]
        kc, vc = hp.impl._views(st["kv"])
        sc = hp.layer._k_scale.view(1) if hp.fp8 else None
        tri = tr.TritonReference(B, hp.hq, D, hp.device)
        o = torch.empty_like(hp.out)

        def fn():
            tri(st["q"], kc, vc, o, st["cu"], 1, st["seq_lens"], w["ctx"], D ** -0.5, st["block_table"], k_scale=sc, v_scale=sc)
        ms, n = kbench.time_fn(fn, min_ms=budget_ms)
        hp.attention()
        torch.cuda.synchronize(hp.device)
        err = float((o.float() - hp.out.float()).abs().max())
        out["triton_unified_attention"] = {"ms": round(ms, 5), "gbs": round(hp.dec_bytes / ms / 1e6, 1), "iters": n,
                                           "source": tr.SOURCE, "max_abs_diff_vs_ours": round(err, 5),
                                           "what": "kernel_unified_attention_3d + reduce_segments, the reference launcher's 16 segments"}
    except Exception as e:  # a comparator must never take the bench line down
        out["triton_unified_attention"] = {"unavailable": f"{type(e).__name__}: {str(e)[:160]}"}
    return out

def run_ours(args):
    from vllm_triton_backend_b200 import _cabi

    rank, world, local = dist_env()
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torchrun for --gpus > 1")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=device)
    # N = 1: C2, the configuration BASELINE.json's metric is quoted on.  N > 1: C4, the north_star's multi-GPU
    # configuration - its 8 KV heads split over the ranks the way vLLM tensor parallelism splits them
    # (vllm/config/model.py:1348-1363), total work fixed ("strong").  --workload overrides.
    name = args.workload or ("c2" if world == 1 else "c4")
    w = WORKLOADS[name]
    if name == "c4":
        assert w["hkv"] % world == 0
        hq, hkv, scaling = w["hq"] // world, w["hkv"] // world, "strong"
    else:
        hq, hkv, scaling = w["hq"], w["hkv"], "weak"
    fp8 = w["kv"] == "fp8"
    B, D = w["batch"], w["d"]
    stream = torch.cuda.current_stream(device)

    def barrier():
        if world > 1:
            import torch.distributed as dist

            dist.barrier()
        torch.cuda.synchronize(device)

    hp = HotPath(w, hq, hkv, device, seed=rank)
    W = max(args.warmup, 3)
    K = args.steps

    # ---- device-resident timing (value): nothing but the step's launches between the two events -----------
    for _ in range(W):
        hp.step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.25)
    _cabi.reset_launch_count()
    t_begin = sampler.mark()
    total_ms = time_steps(hp, K, 0, barrier, stream)
    t_end = sampler.mark()
    launches = _cabi.launch_count()
    dominant = _cabi.last_kernel()

    # ---- roofline of the dominant kernel: its own pass, every launch bracketed -----------------------------
    att_ms = roofline_pass(hp, max(32, args.roofline_launches), stream)

    # ---- end to end through the plugin with HOST buffers (e2e) ---------------------------------------------
    e2e_ms, h2d, d2h, depth = e2e_pass(hp, K, W, barrier, args.e2e_serial)
    clocks = sampler.stop(t_begin, t_end)

    # ---- context (rank 0): the reference's kernels on the same tensors; the other decode configuration -------
    comparators = None
    if rank == 0 and not args.no_comparators:
        comparators = same_gpu_comparators(hp)
    dec_bytes, step_bytes = hp.dec_bytes, hp.step_bytes
    hp.free()
    del hp
    other = {}
    if not args.no_extras:
        # N = 1: the C4 decode on this one GPU (all 8 KV heads) and its 8-GPU shard shape (1 KV head) - the strong-
        # scaling reference points.  N > 1: one C2-shaped replica per rank (round 1's "weak" number).
        todo = [("c4", WORKLOADS["c4"]["hq"], WORKLOADS["c4"]["hkv"]), ("c4", 8, 1)] if world == 1 and name == "c2" else \
               ([("c2", WORKLOADS["c2"]["hq"], WORKLOADS["c2"]["hkv"])] if name == "c4" else [])
        for wl, ohq, ohkv in todo:
            o = HotPath(WORKLOADS[wl], ohq, ohkv, device, seed=rank)
            k2 = max(20, min(K, 200))
            ms = time_steps(o, k2, W, barrier, stream) / k2
            t = torch.tensor([ms], dtype=torch.float64, device=device)
            if world > 1:
                import torch.distributed as dist

                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.cpu())
            mult = world if wl == "c2" else 1
            other[f"{wl}_{ohq}q{ohkv}kv_per_gpu"] = {
                "ms_per_step": round(ms, 5), "gbs": round(mult * o.step_bytes / ms / 1e6, 1), "n_gpus": world,
                "what": ("one replica per rank, sum over ranks" if wl == "c2" else
                         ("every KV head on one GPU" if ohkv == 8 else "the shard one of 8 GPUs holds (1 KV head)"))}
            o.free()
            del o

    pre = None
    if not args.no_prefill:
        torch.cuda.empty_cache()
        pre = measure_prefill(device, world)

    # ---- reduce over ranks ------------------------------------------------------------------------------------
    att_sorted = sorted(att_ms)
    att_med = att_sorted[len(att_sorted) // 2]
    stats = torch.tensor([total_ms, e2e_ms, att_med], dtype=torch.float64, device=device)
    if world > 1:
        import torch.distributed as dist

        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms, att_med = (float(x) for x in stats.cpu())

    # ---- CPU baseline (rank 0, N=1 only) ------------------------------------------------------------------------
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_baseline = run_cpu_reference(w, hq, hkv, budget_s=args.cpu_seconds, seed=0)

    if rank == 0:
        peaks = measured_peaks()
        ms_per_step = total_ms / K
        gbs = world * step_bytes / (ms_per_step * 1e-3) / 1e9
        e2e_gbs = world * step_bytes / (e2e_ms / K * 1e-3) / 1e9
        achieved = dec_bytes / (att_med * 1e-3) / 1e9
        p20, p50, p80 = _percentiles(att_ms, (20, 50, 80))
        line = {
            "metric": "decode_kv_gbps", "value": round(gbs, 2), "unit": "GB/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": round(ms_per_step, 5), "higher_is_better": True,
            "scaling": scaling, "vs_baseline": None, "dtype": "bf16" if not fp8 else "bf16 (fp8-e4m3 KV)",
            "data": "synthetic",
            "config": {"workload": w["desc"], "batch": B, "ctx": w["ctx"], "q_heads_per_gpu": hq, "kv_heads_per_gpu": hkv,
                       "head_dim": D, "block_size": w["bs"], "kv_cache_dtype": w["kv"],
                       "step": "reshape_and_cache_flash(64 new tokens) + decode attention (split-KV, fused merge); the split-KV "
                               "work plan is built once per step's metadata, as the metadata builder does for all layers",
                       "l2": f"KV working set {dec_bytes / 1e9:.2f} GB per GPU >> 126 MB L2; no flush needed",
                       "parallelism": f"kv-head shard x{world} (no collective)"},
            "clocks": clocks,
            "e2e": {"value": round(e2e_gbs, 2), "unit": "GB/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": round(e2e_ms / K, 5),
                    "api": "StepPipeline(depth=%d): one pinned arena per step (q/k/v/metadata) uploaded on a copy stream, ops.DecodePlan.build + B200AttentionImpl.do_kv_cache_update + .forward, result read back on a second copy stream; the host reads every step's result when it reuses the slot" % depth},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": round(achieved, 2), "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": round(achieved / peaks["hbm_gbs"], 4),
                         "traffic": args.ncu_traffic if args.ncu_traffic is not None else (committed_ncu_traffic() if name == "c2" else None),
                         "kernel": "attn_decode_kernel (split-KV, merge of partials fused in)", "peak_source": peaks["source"],
                         "frac_of_nominal_8TBs": round(achieved / 8000.0, 4), "algorithmic_bytes_per_launch": dec_bytes,
                         "avg_launch_ms": round(att_med, 5), "launch_ms_p20_p50_p80": [round(p20, 5), round(p50, 5), round(p80, 5)],
                         "launch_ms_mean": round(statistics.mean(att_ms), 5), "launches_timed": len(att_ms),
                         "timing": "median of CUDA-event pairs on the launching stream around every attention launch of a "
                                   "separate pass (max over ranks); the timed region of `value` carries no events"},
            "extras": {"decode_tokens_per_s": round(world * B / (ms_per_step * 1e-3), 1),
                       "e2e_tokens_per_s": round(world * B / (e2e_ms / K * 1e-3), 1), "last_kernel": dominant,
                       "other_decode_configs": other},
        }
        if comparators is not None:
            line["same_gpu_comparators"] = comparators
        if pre is not None:
            # per-rank numbers of rank 0 (every rank runs the same shapes); whole-job = x world
            line["prefill"] = {
                "bound": "tensor", "unit": "TFLOP/s", "peak": peaks["bf16_tflops"], "peak_source": peaks["source"],
                "full_8k_prompt": {"tflops_per_gpu": round(pre["full8k"]["tflops"], 1), "ms": round(pre["full8k"]["ms"], 4),
                                   "frac": round(pre["full8k"]["tflops"] / peaks["bf16_tflops"], 4),
                                   "tokens_per_s": round(world * pre["full8k"]["tokens"] / pre["full8k"]["ms"] * 1e3, 1)},
                "c3_chunked_prefill_mixed": {"tflops_per_gpu": round(pre["mixed"]["tflops"], 1), "ms": round(pre["mixed"]["ms"], 4),
                                             "frac": round(pre["mixed"]["tflops"] / peaks["bf16_tflops"], 4),
                                             "tokens_per_s": round(world * pre["mixed"]["tokens"] / pre["mixed"]["ms"] * 1e3, 1),
                                             **_mixed_bounds(pre["mixed"], peaks)},
                "roofline": {"bound": "tensor", "achieved": round(pre["full8k"]["tflops"], 1), "peak": peaks["bf16_tflops"],
                             "unit": "TFLOP/s", "frac": round(pre["full8k"]["tflops"] / peaks["bf16_tflops"], 4),
                             "kernel": "attn_prefill_tc_kernel", "algorithmic_flops_per_launch": pre["full8k"]["flops"],
                             "avg_launch_ms": round(pre["full8k"]["ms"], 5), "traffic": None,
                             "ncu": "profiles/: sm__pipe_tensor* of the committed --set full capture of this kernel"},
                "kernel": "attn_prefill_tc_kernel (tcgen05/TMEM) + short/decode rows via attn_prefill_kernel / attn_decode_kernel",
            }
        if cpu_baseline is not None:
            line["cpu_baseline"] = cpu_baseline
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist

        dist.destroy_process_group()


def run_cpu_reference(w, hq, hkv, budget_s, seed, state=None):
    """The oracle (CPU restatement of the reference path: cache write + paged decode
    attention) on a bounded sample of the same workload, all host threads."""
    # the one place bench.py executes oracle/: the CPU baseline / reference arm
    from oracle.cache_write import reshape_and_cache_flash_ref
    from oracle.paged_attention import unified_attention_ref
    from vllm_triton_backend_b200.roofline import decode_bytes

    # torchrun exports OMP_NUM_THREADS=1 to its workers: a CPU arm timed that way is a 1-thread arm
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    B, ctx, D, bs = w["batch"], w["ctx"], w["d"], w["bs"]
    fp8 = w["kv"] == "fp8"
    sample_b = min(B, 8 if ctx <= 8192 else 2)  # sequences per CPU pass
    pages = ctx // bs
    g = torch.Generator().manual_seed(seed)
    NB = sample_b * pages + 8
    if fp8:
        kc = ((torch.rand(NB, bs, hkv, D, generator=g) * 2 - 1) / 0.5).to(torch.float8_e4m3fn)
        vc = ((torch.rand(NB, bs, hkv, D, generator=g) * 2 - 1) / 0.5).to(torch.float8_e4m3fn)
    else:
        kc = (torch.rand(NB, bs, hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
        vc = (torch.rand(NB, bs, hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
    bt = torch.randperm(NB, generator=g)[: sample_b * pages].to(torch.int32).view(sample_b, pages)
    q = (torch.rand(sample_b, hq, D, generator=g) * 2 - 1).to(torch.bfloat16)
    k_new = (torch.rand(sample_b, hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
    v_new = (torch.rand(sample_b, hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
    slots = bt[:, -1].to(torch.int64) * bs + (ctx - 1) % bs
    cu = list(range(sample_b + 1))
    sl = [ctx] * sample_b
    sc = 0.5 if fp8 else 1.0

    def one_pass():
        reshape_and_cache_flash_ref(k_new, v_new, kc, vc, slots, "fp8" if fp8 else "auto", sc, sc)
        unified_attention_ref(q, kc, vc, cu, sl, bt, D ** -0.5, k_scale=sc, v_scale=sc)

    one_pass()  # warm
    n = 0
    t0 = time.perf_counter()
    while True:
        one_pass()
        n += 1
        el = time.perf_counter() - t0
        if el >= budget_s or n >= 1000:
            break
    kvb = 1 if fp8 else 2
    nbytes = decode_bytes(sl, hq, hkv, D, bs, kvb)
    return {"value": round(n * nbytes / el / 1e9, 4), "unit": "GB/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{sample_b} of {B} sequences (ctx {ctx}) per pass, {n} passes in {el:.1f} s, torch fp32 CPU oracle "
                      f"(cache write + paged decode attention)",
            "tokens_per_s": round(n * sample_b / el, 2)}


def run_triton_interpret_c1(timeout_s=150):
    """SURVEY.md 8(d)'s other CPU row: the reference's own Triton kernels under TRITON_INTERPRET=1 on C1 (batch 4,
    32q/8kv, D=128, ctx 512, block 16, fp32), one call, in a subprocess (the interpreter is chosen when triton is
    imported).  Needs the reference kernel file: /root/reference here, baseline/_ref (put there by build()) on
    the GPU box."""
    code = (
        "import os,sys,time,json,torch\n"
        "sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "from oracle import reference_runner as rr\n"
        "from helpers import make_paged_inputs\n"
        "if not rr.reference_available(): print(json.dumps({'unavailable': 'reference kernel file not present'})); sys.exit(0)\n"
        "mod = rr.load_reference_kernels(16, 32)\n"
        "inp = make_paged_inputs([(1,512)]*4, 32, 8, 128, 16, dtype=torch.float32, seed=0)\n"
        "out = torch.empty_like(inp['q'])\n"
        "one = torch.ones(1)\n"
        "t0 = time.perf_counter()\n"
        "mod.unified_attention(q=inp['q'], k=inp['k_cache'], v=inp['v_cache'], out=out, cu_seqlens_q=inp['cu_seqlens_q'], max_seqlen_q=1, seqused_k=inp['seq_lens'], max_seqlen_k=512, avg_seqlen_q=1, avg_seqlen_k=512, softmax_scale=128**-0.5, causal=True, window_size=(-1,-1), block_table=inp['block_table'], softcap=0, q_descale=None, k_descale=one, v_descale=one)\n"
        "el = time.perf_counter() - t0\n"
        "print(json.dumps({'seconds_per_call': round(el, 3), 'decode_tokens_per_s': round(4 / el, 4), 'cores': os.cpu_count(), 'config': 'C1: batch 4, 32q/8kv, D=128, ctx 512, block 16, fp32', 'kernels': 'kernel_unified_attention_3d + reduce_segments of the reference file, Triton CPU interpreter'}))\n"
    ) % (ROOT, os.path.join(ROOT, "tests"))
    env = dict(os.environ, TRITON_INTERPRET="1", CUDA_VISIBLE_DEVICES="")
    env.pop("OMP_NUM_THREADS", None)
    try:
        r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=timeout_s, cwd="/tmp")
        lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
        return json.loads(lines[-1]) if lines else {"unavailable": (r.stderr or "no output")[-200:]}
    except Exception as e:
        return {"unavailable": f"{type(e).__name__}: {str(e)[:160]}"}


def run_reference_arm(args):
    rank, world, local = dist_env()
    if rank != 0:
        return
    args.workload = args.workload or ("c2" if world <= 1 else "c4")
    w = WORKLOADS[args.workload]
    hq, hkv = w["hq"], w["hkv"]
    if args.workload == "c4":
        hq, hkv = hq // max(world, 1), hkv // max(world, 1)
    K = max(args.steps, 1)
    per_step = max(2.0, min(20.0, 60.0 / K))
    for _ in range(min(args.warmup, 1)):
        run_cpu_reference(w, hq, hkv, budget_s=0.5, seed=0)
    vals = []
    t0 = time.perf_counter()
    last = None
    for _ in range(K):
        last = run_cpu_reference(w, hq, hkv, budget_s=per_step, seed=0)
        vals.append(last["value"])
        if time.perf_counter() - t0 > 150:
            break
    v = statistics.mean(vals)
    fp8 = w["kv"] == "fp8"
    cb = dict(last)
    cb["value"] = round(v, 4)
    line = {
        "impl": "reference", "metric": "decode_kv_gbps", "value": round(v, 4), "unit": "GB/s", "n_gpus": max(world, 1),
        "steps": len(vals), "warmup": min(args.warmup, 1), "ms_per_step": None, "higher_is_better": True,
        "scaling": "strong" if args.workload == "c4" else "weak", "vs_baseline": None,
        "dtype": "bf16" if not fp8 else "bf16 (fp8-e4m3 KV)", "data": "synthetic",
        "config": {"workload": w["desc"], "batch": w["batch"], "ctx": w["ctx"], "q_heads_per_gpu": hq,
                   "kv_heads_per_gpu": hkv, "head_dim": w["d"], "block_size": w["bs"], "kv_cache_dtype": w["kv"],
                   "note": "reference arm = CPU restatement of the reference path (oracle/), host cores only; "
                           "each step is a bounded sample of the batch"},
        "cpu_baseline": cb,
        "e2e": {"value": round(v, 4), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    if not args.no_extras:
        line["cpu_baseline"]["triton_interpret_c1"] = run_triton_interpret_c1()
    print(json.dumps(line), flush=True)


def _percentiles(xs, ps):
    ys = sorted(xs)
    return [ys[min(len(ys) - 1, max(0, int(round(p / 100.0 * (len(ys) - 1)))))] for p in ps]


def committed_ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the decode kernel on
    this workload, from the committed `ncu --set full` capture (profiles/ncu_traffic.json)."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return int(json.load(f)["attn_decode_kernel_c2"]["dram_bytes_per_launch"])
    except (OSError, KeyError, ValueError):
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS),
                    help="default: c2 at --gpus 1, c4 (KV heads sharded over the ranks, strong scaling) at --gpus > 1")
    ap.add_argument("--roofline-launches", type=int, default=64, help="bracketed attention launches of the roofline pass (>= 32)")
    ap.add_argument("--no-comparators", action="store_true", help="skip the same-GPU Triton reference timing")
    ap.add_argument("--no-extras", action="store_true", help="skip the other decode configuration (C4 at N=1, C2 replicas at N>1)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-serial", action="store_true", help="e2e without overlap: upload, compute, read back, sync, every step")
    ap.add_argument("--no-prefill", action="store_true", help="skip the prefill TFLOP/s section")
    ap.add_argument("--ncu-traffic", type=float, default=None,
                    help="dram bytes per launch of the decode kernel from the committed ncu capture (profiles/)")
    ap.add_argument("--tune", type=int, default=0,
                    help="A/B switch word for the library (b200attn_set_tuning `reserved`; 512 = no programmatic dependent launch)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        if args.tune:
            from vllm_triton_backend_b200 import _cabi

            _cabi.set_tuning(0, 0, 0, args.tune)
        run_ours(args)


if __name__ == "__main__":
    main()
