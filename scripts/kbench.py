#!/usr/bin/env python
"""Kernel micro-benchmarks for tuning (not the driver's bench.py contract).

    python scripts/kbench.py decode  [--workload c2|c4|c2var] [--hpc H] [--stages S] [--split X] [--hkv N --hq N]
    python scripts/kbench.py prefill [--shape full8k|chunk|mixed|full2k] [--force 2|4]
    python scripts/kbench.py cache   [--tokens T] [--fp8]
    python scripts/kbench.py sweep   (decode knob sweep on c2 and c4)

Every line printed is JSON: time per call (CUDA events around >= 50 ms of
back-to-back launches after warm-up; working sets are >> L2 for decode, and an
L2 flush variant is available with --flush), algorithmic GB/s or TFLOP/s and the
fraction of the measured peak (MEASURED_PEAKS.json).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from vllm_triton_backend_b200 import _cabi, ops  # noqa: E402
from vllm_triton_backend_b200.roofline import attention_flops, cache_write_bytes, decode_bytes  # noqa: E402


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d["bf16_tflops"])
    return 6650.0, 1590.0


LAST_CLOCK = None


def sample_clock():
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(0)
        return dict(sm_mhz=pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM),
                    power_w=round(pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0, 1))
    except Exception:
        return None


def time_fn(fn, min_ms=60.0, warmup=5, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # calibrate
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    one = max(e0.elapsed_time(e1), 1e-3)
    n = max(5, min(2000, int(min_ms / one)))
    if flush is None:
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        global LAST_CLOCK
        LAST_CLOCK = sample_clock()  # launches are still running: the clock under load
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n, n
    tot = 0.0
    n = min(n, 50)
    for _ in range(n):
        flush.zero_()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / n, n


def triton_reference():
    """tests/triton_reference.py: the reference Triton kernels as shipped by the installed vLLM."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import triton_reference as tr

    why = tr.available()
    if why:
        raise SystemExit(f"--impl triton unavailable: {why}")
    return tr


def kernel_name(args):
    return "triton kernel_unified_attention (vLLM %s)" % __import__("vllm").__version__ if getattr(args, "impl", "ours") == "triton" else _cabi.last_kernel()


def graphed(fn, inner=20):
    """`inner` back-to-back calls captured in one CUDA graph (the do_bench_cudagraph-style variant of
    SURVEY.md §8d): returns a callable replaying the graph and the number of calls per replay."""
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(inner):
            fn()
    return g.replay, inner


def make_kv(NB, bs, hkv, D, fp8, device, g, layout="planes"):
    shape = (2, NB, bs, hkv, D) if layout == "planes" else (NB, 2, bs, hkv, D)
    dt = torch.uint8 if fp8 else torch.bfloat16
    kv = torch.empty(shape, dtype=dt, device=device)
    flat = kv.view(-1)
    chunk = 1 << 26
    for o in range(0, flat.numel(), chunk):
        n = min(chunk, flat.numel() - o)
        x = torch.rand(n, device=device, generator=g) * 2 - 1
        if fp8:
            flat[o:o + n] = (x / 0.5).to(torch.float8_e4m3fn).view(torch.uint8)
        else:
            flat[o:o + n] = x.to(torch.bfloat16)
    if fp8:
        kv = kv.view(torch.float8_e4m3fn)
    return kv.unbind(0) if layout == "planes" else kv.unbind(1)


def decode_case(args, device):
    g = torch.Generator(device=device).manual_seed(0)
    if args.workload == "c4":
        B, ctx, hq, hkv, D, bs, fp8 = 64, 32768, 64, 8, 128, 16, True
    else:
        B, ctx, hq, hkv, D, bs, fp8 = 64, 4096, 32, 8, 128, 16, False
    if args.hq:
        hq = args.hq
    if args.hkv:
        hkv = args.hkv
    if args.batch:
        B = args.batch
    if args.ctx:
        ctx = args.ctx
    if args.bs:
        bs = args.bs
    if args.fp8 or args.pth:
        fp8 = True
    if args.workload == "c2var":
        pat = [0.1, 0.4, 0.5, 1.0, 0.2]  # reference scripts/benchmark.py:115
        lens = [max(1, int(ctx * pat[i % len(pat)])) for i in range(B)]
    else:
        lens = [ctx] * B
    pages = [(n + bs - 1) // bs for n in lens]
    NB = sum(pages) + 1024
    kc, vc = make_kv(NB, bs, hkv, D, fp8, device, g, args.layout)
    perm = torch.randperm(NB, device=device, generator=g).to(torch.int32)
    bt = torch.zeros(B, max(pages), dtype=torch.int32, device=device)
    o = 0
    for i, n in enumerate(pages):
        bt[i, :n] = perm[o:o + n]
        o += n
    sl = torch.tensor(lens, dtype=torch.int32, device=device)
    cu = torch.arange(B + 1, dtype=torch.int32, device=device)
    q = (torch.rand(B, hq, D, device=device, generator=g) * 2 - 1).to(torch.bfloat16)
    out = torch.empty_like(q)
    ws = torch.zeros(ops.workspace_bytes(B, B, hq, hkv, D), dtype=torch.uint8, device=device)
    sc = torch.tensor([0.5], dtype=torch.float32, device=device) if fp8 else None
    nbytes = decode_bytes(lens, hq, hkv, D, bs, 1 if fp8 else 2)
    pth_kw = {}
    if args.pth:
        # scale caches [NB, bs, Hkv] laid out [Hkv][bs] inside a block, as in the plugin's allocation
        # (ops.per_token_head_cache_views); their bytes count as traffic too
        if args.pth == "int8":
            kc, vc = kc.view(torch.int8), vc.view(torch.int8)
        ksc = (torch.rand(NB, hkv, bs, device=device, generator=g) * 0.01 + 0.001).transpose(1, 2)
        vsc = (torch.rand(NB, hkv, bs, device=device, generator=g) * 0.01 + 0.001).transpose(1, 2)
        pth_kw = dict(kv_quant_mode=2 if args.pth == "int8" else 3, k_scale_cache=ksc, v_scale_cache=vsc)
        sc = None
        nbytes += sum(lens) * hkv * 2 * 4
    plan = None
    if getattr(args, "plan", False):
        # what the metadata builder does once per step: the split-KV work plan, shared by every layer's launch
        _cabi.set_tuning(args.hpc, args.ctas, args.stages, args.split)
        plan = ops.DecodePlan(device, B, hq, hkv, D, bs, torch.bfloat16, torch.float8_e4m3fn if fp8 else torch.bfloat16)
        plan.build(cu, sl, 1)

    def fn():
        ops.unified_attention(q=q, k=kc, v=vc, out=out, cu_seqlens_q=cu, max_seqlen_q=1, seqused_k=sl,
                              max_seqlen_k=max(lens), softmax_scale=D ** -0.5, causal=True, window_size=(-1, -1),
                              block_table=bt, softcap=0, q_descale=None, k_descale=sc, v_descale=sc, workspace=ws,
                              force_selection=args.force, decode_plan=plan, **pth_kw)
    if getattr(args, "impl", "ours") == "triton":
        tri = triton_reference().TritonReference(B, hq, D, device)

        def fn():  # noqa: F811 - the reference kernels on the same tensors
            tri(q, kc, vc, out, cu, 1, sl, max(lens), D ** -0.5, bt, k_scale=sc, v_scale=sc, use_3d=args.force != 2,
                k_scale_cache=pth_kw.get("k_scale_cache"), v_scale_cache=pth_kw.get("v_scale_cache"))
    return fn, nbytes, dict(B=B, ctx=ctx, hq=hq, hkv=hkv, D=D, bs=bs, fp8=fp8, workload=args.workload, layout=args.layout,
                            plan=plan is not None, per_token_head=args.pth)


def decode_trace(fn, device):
    """Per-CTA phase times of one decode launch (b200attn_debug_set_decode_trace)."""
    lib = _cabi.load_library()
    nsm = torch.cuda.get_device_properties(device).multi_processor_count
    buf = torch.zeros(nsm * 16, dtype=torch.int64, device=device)
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    lib.b200attn_debug_set_decode_trace(buf.data_ptr())
    fn()
    torch.cuda.synchronize()
    lib.b200attn_debug_set_decode_trace(None)
    t = buf.cpu().view(nsm, 16).double()
    t = t[t[:, 0] > 0]
    span_ns = float(t[:, 6].max() - t[:, 0].min())
    ghz = float(((t[:, 5] - t[:, 1]) / (t[:, 6] - t[:, 0]).clamp(min=1)).median())  # cycles per ns
    us = lambda cyc: cyc / ghz / 1e3

    def stat(x):
        return dict(mean=round(float(x.mean()), 2), min=round(float(x.min()), 2), max=round(float(x.max()), 2))
    return dict(kind="decode_trace", ctas=int(t.shape[0]), kernel_span_us=round(span_ns / 1e3, 2), sm_ghz=round(ghz, 3),
                start_skew_us=stat((t[:, 0] - t[:, 0].min()) / 1e3), first_run_us=stat(us(t[:, 2] - t[:, 1])),
                first_data_us=stat(us(t[:, 3] - t[:, 2])), streaming_us=stat(us(t[:, 4] - t[:, 3])),
                tail_us=stat(us(t[:, 5] - t[:, 4])), cta_total_us=stat((t[:, 6] - t[:, 0]) / 1e3),
                end_skew_us=stat((t[:, 6].max() - t[:, 6]) / 1e3), runs_per_cta=stat(t[:, 7]),
                finalize_us=stat(us(t[:, 8])), consumer_wait_full_us=stat(us(t[:, 9])), producer_wait_empty_us=stat(us(t[:, 10])),
                producer_switch_us=stat(us(t[:, 11])))


def run_decode(args, device):
    hbm, _ = peaks()
    fn, nbytes, cfg = decode_case(args, device)
    _cabi.set_tuning(args.hpc, args.ctas, args.stages, args.split)
    if args.trace:
        print(json.dumps(dict(**decode_trace(fn, device), **cfg)), flush=True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device) if args.flush else None
    inner = 1
    if args.graph:
        fn, inner = graphed(fn)
    ms, n = time_fn(fn, flush=flush)
    ms /= inner
    cfg["cuda_graph"] = bool(args.graph)
    gbs = nbytes / ms / 1e6
    print(json.dumps(dict(kind="decode", **cfg, hpc=args.hpc, stages=args.stages, split=args.split, ctas=args.ctas,
                          ms=round(ms, 5), gbs=round(gbs, 1), frac_measured=round(gbs / hbm, 4),
                          frac_8tbs=round(gbs / 8000, 4), iters=n, kernel=kernel_name(args), flush=bool(args.flush),
                          impl=args.impl)),
          flush=True)
    _cabi.set_tuning(0, 0, 0, 0)


PREFILL_SHAPES = {
    # list of (q_len, kv_len)
    "full8k": [(8192, 8192)],
    "full2k": [(2048, 2048)] * 4,
    "full1k": [(1024, 1024)] * 8,
    "chunk": [(2048, 6144), (2048, 8192)],
    # the multi-token sequences of C3 alone (what the tensor-bound kernels do in the mixed batch)
    "mixed_prompts": None,
    # C3: 32 decode rows + fresh prompt + 2 chunked continuations + 4 prefix hits
    "mixed": None,
}


def mixed_c3(seed=0):
    import random

    rnd = random.Random(seed)
    dec = [(1, rnd.randint(128, 8192)) for _ in range(32)]
    big = [(2048, 2048), (2048, 6144), (1024, 7168)]
    hits = [(rnd.randint(128, 1024), 0) for _ in range(4)]
    hits = [(q, q + (1024 if i % 2 == 0 else 2048)) for i, (q, _) in enumerate(hits)]
    seqs = []
    rest = big + hits
    # alternate decode / prefill as the reference harness does (scripts/benchmark.py:1101-1112)
    while dec or rest:
        if rest:
            seqs.append(rest.pop(0))
        if dec:
            seqs.append(dec.pop(0))
    return seqs


def prefill_case(args, device):
    g = torch.Generator(device=device).manual_seed(0)
    hq, hkv, D, bs = args.hq or 32, args.hkv or 8, getattr(args, "d", 0) or 128, args.bs or 16
    lens = PREFILL_SHAPES[args.shape] or mixed_c3()
    if args.shape == "mixed_prompts":
        lens = [x for x in lens if x[0] > 1]
    q_lens = [a for a, _ in lens]
    kv_lens = [b for _, b in lens]
    pages = [(n + bs - 1) // bs for n in kv_lens]
    NB = sum(pages) + 64
    kc, vc = make_kv(NB, bs, hkv, D, args.fp8, device, g, args.layout)
    perm = torch.randperm(NB, device=device, generator=g).to(torch.int32)
    bt = torch.zeros(len(lens), max(pages), dtype=torch.int32, device=device)
    o = 0
    for i, n in enumerate(pages):
        bt[i, :n] = perm[o:o + n]
        o += n
    T = sum(q_lens)
    q = (torch.rand(T, hq, D, device=device, generator=g) * 2 - 1).to(torch.bfloat16)
    out = torch.empty_like(q)
    cu = torch.tensor([0] + list(torch.tensor(q_lens).cumsum(0)), dtype=torch.int32, device=device)
    sl = torch.tensor(kv_lens, dtype=torch.int32, device=device)
    sc = torch.tensor([0.5], dtype=torch.float32, device=device) if args.fp8 else None
    ws_bytes = ops.workspace_bytes(T, len(lens), hq, hkv, D)
    if args.fp8 and not getattr(args, "no_staging", False):  # room to stage the prefill sequences (csrc/kv_stage.cu)
        ws_bytes += ops.fp8_staging_bytes(len(lens), sum(kv for q, kv in lens if q > 16), bs, hkv, D)
    ws = torch.zeros(ws_bytes, dtype=torch.uint8, device=device)
    flops = attention_flops(q_lens, kv_lens, hq, D)
    softcap = float(getattr(args, "softcap", 0.0) or 0.0)

    def fn():
        ops.unified_attention(q=q, k=kc, v=vc, out=out, cu_seqlens_q=cu, max_seqlen_q=max(q_lens), seqused_k=sl,
                              max_seqlen_k=max(kv_lens), softmax_scale=D ** -0.5, causal=True, window_size=(-1, -1),
                              block_table=bt, softcap=softcap, q_descale=None, k_descale=sc, v_descale=sc, workspace=ws,
                              force_selection=args.force)
    if getattr(args, "impl", "ours") == "triton":
        tri = triton_reference().TritonReference(len(lens), hq, D, device)

        def fn():  # noqa: F811
            tri(q, kc, vc, out, cu, max(q_lens), sl, max(kv_lens), D ** -0.5, bt, k_scale=sc, v_scale=sc, softcap=softcap)
    # bytes of the KV the single-token (decode) rows stream: the HBM-bound share of a mixed batch
    dec_bytes = sum(kv for ql, kv in lens if ql == 1) * hkv * D * 2 * (1 if args.fp8 else 2)
    return fn, flops, dict(shape=args.shape, T=T, seqs=len(lens), hq=hq, hkv=hkv, D=D, bs=bs, fp8=bool(args.fp8),
                           softcap=softcap, decode_rows=sum(1 for ql, _ in lens if ql == 1), decode_row_kv_bytes=dec_bytes)


def run_prefill(args, device):
    _, tf = peaks()
    fn, flops, cfg = prefill_case(args, device)
    _cabi.set_tuning(args.hpc, 0, 0, args.tune)
    cfg["tune"], cfg["hpc"] = args.tune, args.hpc
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device) if args.flush else None
    ms, n = time_fn(fn, flush=flush)
    tfs = flops / ms / 1e9
    print(json.dumps(dict(kind="prefill", **cfg, force=args.force, ms=round(ms, 5), tflops=round(tfs, 1),
                          frac_measured=round(tfs / tf, 4), tokens_per_s=round(cfg["T"] / ms * 1e3, 1), iters=n,
                          kernel=kernel_name(args), clock=LAST_CLOCK, impl=args.impl)), flush=True)


def run_cache(args, device):
    hbm, _ = peaks()
    g = torch.Generator(device=device).manual_seed(0)
    T, hkv, D, bs = args.tokens, args.hkv or 8, 128, args.bs or 16
    NB = (T + bs - 1) // bs + 64
    kc, vc = make_kv(NB, bs, hkv, D, args.fp8, device, g, args.layout)
    key = (torch.rand(T, hkv, D, device=device, generator=g) * 2 - 1).to(torch.bfloat16)
    val = (torch.rand(T, hkv, D, device=device, generator=g) * 2 - 1).to(torch.bfloat16)
    slots = torch.randperm(NB * bs, device=device, generator=g)[:T].to(torch.int64)
    sc = torch.tensor([0.5], dtype=torch.float32, device=device)
    nbytes = cache_write_bytes(T, hkv, D, 2, 1 if args.fp8 else 2)

    def fn():
        ops.reshape_and_cache_flash(key, val, kc, vc, slots, "fp8" if args.fp8 else "auto", sc, sc)
    if args.pth:
        kc, vc = make_kv(NB, bs, hkv, D, True, device, g, args.layout)
        if args.pth == "int8":
            kc, vc = kc.view(torch.int8), vc.view(torch.int8)
        ksc = torch.ones(NB, hkv, bs, device=device).transpose(1, 2)
        vsc = torch.ones(NB, hkv, bs, device=device).transpose(1, 2)
        nbytes = cache_write_bytes(T, hkv, D, 2, 1) + T * hkv * 2 * 4

        def fn():  # noqa: F811
            ops.reshape_and_cache_flash_per_token_head(key, val, kc, vc, ksc, vsc, slots)
        if args.impl == "triton":
            from vllm.v1.attention.ops.triton_reshape_and_cache_flash import (
                triton_reshape_and_cache_flash_per_token_head_quant as theirs,
            )

            def fn():  # noqa: F811
                theirs(key, val, kc, vc, ksc, vsc, slots)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device) if args.flush else None
    _cabi.set_tuning(0, 0, 0, args.tune)
    ms, n = time_fn(fn, flush=flush)
    _cabi.set_tuning(0, 0, 0, 0)
    gbs = nbytes / ms / 1e6
    print(json.dumps(dict(kind="cache_write", tune=args.tune, tokens=T, hkv=hkv, D=D, bs=bs, fp8=bool(args.fp8), ms=round(ms, 5),
                          gbs=round(gbs, 1), frac_measured=round(gbs / hbm, 4), iters=n, kernel=kernel_name(args),
                          flush=bool(args.flush), per_token_head=args.pth, impl=args.impl)), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("kind", choices=["decode", "prefill", "cache", "sweep"])
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--shape", default="full8k", choices=sorted(PREFILL_SHAPES))
    ap.add_argument("--hpc", type=int, default=0)
    ap.add_argument("--ctas", "--pool", dest="ctas", type=int, default=0,
                    help="decode: share of a CTA's steps that goes to the dynamic pool, in 1/16 (0 = default 3, 100 = none)")
    ap.add_argument("--plan", action="store_true", help="decode: build the split-KV plan once (as the metadata builder does)")
    ap.add_argument("--stages", type=int, default=0)
    ap.add_argument("--split", type=int, default=0)
    ap.add_argument("--hq", type=int, default=0)
    ap.add_argument("--hkv", type=int, default=0)
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--ctx", type=int, default=0)
    ap.add_argument("--bs", type=int, default=0)
    ap.add_argument("--d", type=int, default=0, help="prefill: head size (default 128)")
    ap.add_argument("--tokens", type=int, default=8192)
    ap.add_argument("--fp8", action="store_true")
    ap.add_argument("--pth", default="", choices=["", "fp8", "int8"],
                    help="decode / cache: per-(token, head) quantised cache (one fp32 scale per cached row)")
    ap.add_argument("--flush", action="store_true")
    ap.add_argument("--layout", default="planes", choices=["planes", "interleaved"])
    ap.add_argument("--force", type=int, default=None)
    ap.add_argument("--softcap", type=float, default=0.0, help="prefill: logit soft cap (0 = off)")
    ap.add_argument("--no-staging", action="store_true", help="fp8 prefill: no staging pool (mma.sync kernel only)")
    ap.add_argument("--graph", action="store_true", help="decode: time 20 calls captured in one CUDA graph")
    ap.add_argument("--trace", action="store_true", help="decode: print per-CTA phase times of one launch first")
    ap.add_argument("--impl", default="ours", choices=["ours", "triton"],
                    help="triton = time the reference Triton kernels (vLLM's copy) on the same tensors")
    ap.add_argument("--tune", type=int, default=0, help="raw b200attn_set_tuning `reserved` word (kernel variant bits)")
    args = ap.parse_args()
    device = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    if args.kind == "decode":
        run_decode(args, device)
    elif args.kind == "prefill":
        run_prefill(args, device)
    elif args.kind == "cache":
        run_cache(args, device)
    else:
        hbm, _ = peaks()
        variants = [("c2", {}), ("c2var", {}), ("c4", {}), ("c4", dict(hq=32, hkv=4)), ("c4", dict(hq=16, hkv=2)),
                    ("c4", dict(hq=8, hkv=1)), ("c2", dict(batch=8)), ("c2", dict(batch=1, ctx=32768)),
                    ("c2", dict(batch=256, ctx=1024)), ("c2", dict(bs=128))]
        for wl, over in variants:
            args.workload = wl
            args.hq = args.hkv = args.batch = args.ctx = args.bs = 0
            for k, v in over.items():
                setattr(args, k, v)
            fn, nbytes, cfg = decode_case(args, device)
            for hpc in (0, 1, 2, 4, 8):
                if hpc > cfg["hkv"]:
                    continue
                for stages in (0, 4):
                    try:
                        _cabi.set_tuning(hpc, 0, stages, 0)
                        ms, n = time_fn(fn, min_ms=30.0)
                        gbs = nbytes / ms / 1e6
                        print(json.dumps(dict(kind="decode", workload=wl, **over, hpc=hpc, stages=stages,
                                              ms=round(ms, 5), gbs=round(gbs, 1), frac_measured=round(gbs / hbm, 4))),
                              flush=True)
                    except Exception as e:  # keep sweeping
                        print(json.dumps(dict(kind="decode", workload=wl, **over, hpc=hpc, stages=stages,
                                              error=str(e)[:200])), flush=True)
            _cabi.set_tuning(0, 0, 0, 0)
            del fn
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
