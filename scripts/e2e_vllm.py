#!/usr/bin/env python
"""C5 (BASELINE.json configs[4]): end-to-end vLLM offline inference through the plugin.

granite-3.1-8b architecture, random-init ("dummy") weights, 64 prompts x 512 random
token ids in, 256 tokens out, greedy, ignore_eos — SURVEY.md §8(d) C5.  No network:
the model directory (config.json only) is written here and the tokenizer is skipped.

    python scripts/e2e_vllm.py --backend b200     # this package, via vllm.general_plugins + CUSTOM
    python scripts/e2e_vllm.py --backend triton   # vLLM's TRITON_ATTN = upstream of the reference kernels
    python scripts/e2e_vllm.py --backend flash    # vLLM's default on this GPU, for scale
    python scripts/e2e_vllm.py --compare a.json b.json   # greedy-token agreement of two runs

The plugin is discovered the way an installed wheel would be: a `.dist-info` with
the entry points of setup.py is put on PYTHONPATH (nothing is written into
site-packages), so the engine-core subprocess loads it as well.  One JSON line per run.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

GRANITE_3_1_8B = {
    # ibm-granite/granite-3.1-8b-instruct config.json (architecture only; weights are random)
    "architectures": ["GraniteForCausalLM"], "model_type": "granite", "hidden_size": 4096, "intermediate_size": 12800,
    "num_hidden_layers": 40, "num_attention_heads": 32, "num_key_value_heads": 8, "vocab_size": 49155,
    "max_position_embeddings": 131072, "rms_norm_eps": 1e-5, "rope_theta": 10000000.0, "hidden_act": "silu",
    "attention_bias": False, "mlp_bias": False, "attention_dropout": 0.0, "tie_word_embeddings": True,
    "attention_multiplier": 0.0078125, "embedding_multiplier": 12.0, "logits_scaling": 16.0, "residual_multiplier": 0.22,
    "initializer_range": 0.02, "torch_dtype": "bfloat16", "bos_token_id": 0, "eos_token_id": 0, "pad_token_id": 0,
    "use_cache": True,
}


def write_model_dir(layers: int) -> str:
    d = os.path.join(tempfile.gettempdir(), f"granite31_8b_random_{layers}l")
    os.makedirs(d, exist_ok=True)
    cfg = dict(GRANITE_3_1_8B, num_hidden_layers=layers)
    with open(os.path.join(d, "config.json"), "w") as f:
        json.dump(cfg, f)
    return d


def expose_plugin() -> str:
    """Put `<tmp>/vllm_triton_backend_b200-0.1.0.dist-info/entry_points.txt` (the entry
    points of setup.py) and the repo root on sys.path / PYTHONPATH."""
    site = os.path.join(tempfile.gettempdir(), "b200attn_plugin_site")
    di = os.path.join(site, "vllm_triton_backend_b200-0.1.0.dist-info")
    os.makedirs(di, exist_ok=True)
    with open(os.path.join(di, "METADATA"), "w") as f:
        f.write("Metadata-Version: 2.1\nName: vllm-triton-backend-b200\nVersion: 0.1.0\n")
    with open(os.path.join(di, "entry_points.txt"), "w") as f:
        f.write("[vllm.general_plugins]\nb200_attn_backend = vllm_triton_backend_b200.backend:register_backend\n")
    for p in (site, ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["PYTHONPATH"] = os.pathsep.join([site, ROOT] + [x for x in os.environ.get("PYTHONPATH", "").split(os.pathsep) if x])
    return site


def run(args) -> dict:
    os.environ.setdefault("HF_HUB_OFFLINE", "1")
    os.environ.setdefault("TRANSFORMERS_OFFLINE", "1")
    os.environ.setdefault("VLLM_NO_USAGE_STATS", "1")
    os.environ.setdefault("VLLM_LOGGING_LEVEL", "WARNING")
    if args.tp > 1:
        args.mp = True  # tensor parallelism runs one worker process per GPU
    if not args.mp:
        os.environ["VLLM_ENABLE_V1_MULTIPROCESSING"] = "0"  # engine core in this process: kernel launch counts visible
    if args.backend == "b200":
        expose_plugin()
    import random

    import torch
    from vllm import LLM, SamplingParams

    backend = {"b200": "CUSTOM", "triton": "TRITON_ATTN", "flash": None}[args.backend]
    model_dir = write_model_dir(args.layers)
    kw = dict(model=model_dir, skip_tokenizer_init=True, load_format="dummy", dtype="bfloat16", seed=0,
              max_model_len=args.max_model_len, gpu_memory_utilization=args.gpu_mem, enable_prefix_caching=args.prefix_caching,
              max_num_seqs=args.max_num_seqs or args.num_prompts, max_num_batched_tokens=args.max_batched_tokens, enforce_eager=args.eager,
              kv_cache_dtype=args.kv_cache_dtype, generation_config="vllm")
    if backend is not None:
        kw["attention_backend"] = backend
    if args.tp > 1:
        # the path shards by KV head exactly as vLLM's tensor parallelism does (vllm/config/model.py:1348-1363): every
        # worker process loads the plugin and runs the kernels on its own heads; the only collective is vLLM's
        # all-reduce behind o_proj, outside this path
        kw.update(tensor_parallel_size=args.tp, distributed_executor_backend="mp")
    t0 = time.perf_counter()
    llm = LLM(**kw)
    t_init = time.perf_counter() - t0

    rnd = random.Random(0)
    vocab = GRANITE_3_1_8B["vocab_size"]
    shared = [rnd.randrange(10, vocab) for _ in range(args.shared_prefix)]
    prompts = [{"prompt_token_ids": shared + [rnd.randrange(10, vocab) for _ in range(args.input_len - len(shared))]}
               for _ in range(args.num_prompts)]
    sp = SamplingParams(max_tokens=args.output_len, ignore_eos=True, temperature=0.0, detokenize=False)
    if args.ragged:
        # continuous batching: prompts and generations of different lengths, more requests than batch slots, so
        # the decode batch grows and shrinks from step to step and prefill chunks share steps with decode rows
        prompts = [{"prompt_token_ids": p["prompt_token_ids"][: rnd.randint(32, args.input_len)]} for p in prompts]
        sp = [SamplingParams(max_tokens=rnd.randint(8, args.output_len), ignore_eos=True, temperature=0.0, detokenize=False)
              for _ in prompts]
    # warm-up on the full batch (same prefill/decode batch shapes as the timed run), so that JIT
    # compilation and autotuning of either backend stay out of the timed region
    warm = SamplingParams(max_tokens=16, ignore_eos=True, temperature=0.0, detokenize=False)
    llm.generate(prompts, warm, use_tqdm=False)
    if args.prefix_caching and hasattr(llm, "reset_prefix_cache"):
        llm.reset_prefix_cache()

    launches = None
    if args.backend == "b200" and not args.mp:
        from vllm_triton_backend_b200 import _cabi
        _cabi.load_library().b200attn_reset_launch_count()
    times = []
    for _ in range(args.repeats):
        if args.prefix_caching and hasattr(llm, "reset_prefix_cache"):
            llm.reset_prefix_cache()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        outs = llm.generate(prompts, sp, use_tqdm=False)
        torch.cuda.synchronize()
        times.append(time.perf_counter() - t0)
    dt = sorted(times)[len(times) // 2]  # median
    if args.backend == "b200" and not args.mp:
        launches = int(_cabi.load_library().b200attn_launch_count())
    toks = [list(o.outputs[0].token_ids) for o in outs]
    n_out = sum(len(t) for t in toks)
    selected = None
    try:  # which backend class the layers really got
        from vllm.config import get_layers_from_vllm_config  # noqa: F401
        core = llm.llm_engine.engine_core
        mr = core.engine_core.model_executor.driver_worker.worker.model_runner if hasattr(core, "engine_core") else None
        if mr is not None:
            selected = sorted({type(g.backend).__name__ if not isinstance(g.backend, type) else g.backend.__name__
                               for groups in mr.attn_groups for g in groups})
    except Exception as e:  # introspection only
        selected = f"unknown ({type(e).__name__})"
    res = {
        "metric": "e2e_generated_tokens_per_s", "value": round(n_out / dt, 1), "unit": "tokens/s", "backend": args.backend,
        "attention_backend": backend or "auto", "selected_backend_classes": selected, "gpu_launches_b200attn": launches,
        "seconds": round(dt, 3), "all_seconds": [round(t, 3) for t in times], "init_seconds": round(t_init, 1), "generated_tokens": n_out,
        "prompt_tokens": args.num_prompts * args.input_len,
        "total_tokens_per_s": round((n_out + args.num_prompts * args.input_len) / dt, 1),
        "config": {"model": f"granite-3.1-8b architecture, {args.layers} layers, random-init bf16 (load_format=dummy)",
                   "num_prompts": args.num_prompts, "input_len": args.input_len, "output_len": args.output_len,
                   "prefix_caching": args.prefix_caching, "shared_prefix": args.shared_prefix,
                   "kv_cache_dtype": args.kv_cache_dtype, "enforce_eager": args.eager,
                   "engine_core": "subprocess" if args.mp else "in-process", "tensor_parallel_size": args.tp},
    }
    if args.save:
        with open(args.save, "w") as f:
            json.dump({"result": res, "tokens": toks}, f)
    return res


def compare(a_path: str, b_path: str) -> dict:
    a, b = json.load(open(a_path)), json.load(open(b_path))
    ta, tb = a["tokens"], b["tokens"]
    same_first = sum(x[0] == y[0] for x, y in zip(ta, tb))
    prefix = []
    for x, y in zip(ta, tb):
        n = 0
        while n < min(len(x), len(y)) and x[n] == y[n]:
            n += 1
        prefix.append(n)
    total = sum(min(len(x), len(y)) for x, y in zip(ta, tb))
    return {"compare": [a["result"]["backend"], b["result"]["backend"]], "sequences": len(prefix),
            "first_token_agreement": round(same_first / max(1, len(prefix)), 4),
            "mean_agreeing_prefix_tokens": round(sum(prefix) / max(1, len(prefix)), 1),
            "fully_identical_sequences": sum(p == min(len(x), len(y)) for p, x, y in zip(prefix, ta, tb)),
            "token_agreement_until_first_divergence": round(sum(prefix) / max(1, total), 4),
            "speedup_a_over_b": round(a["result"]["value"] / b["result"]["value"], 4),
            "note": "random-init weights give near-flat logits, so greedy decoding diverges at the first bf16-level tie; "
                    "agreement until that point is the meaningful number"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="b200", choices=["b200", "triton", "flash"])
    ap.add_argument("--layers", type=int, default=40)
    ap.add_argument("--num-prompts", type=int, default=64)
    ap.add_argument("--input-len", type=int, default=512)
    ap.add_argument("--output-len", type=int, default=256)
    ap.add_argument("--shared-prefix", type=int, default=0, help="leading tokens shared by all prompts (prefix-cache hits)")
    ap.add_argument("--max-model-len", type=int, default=1024)
    ap.add_argument("--max-batched-tokens", type=int, default=8192)
    ap.add_argument("--gpu-mem", type=float, default=0.5)
    ap.add_argument("--kv-cache-dtype", default="auto")
    ap.add_argument("--prefix-caching", action="store_true")
    ap.add_argument("--eager", action="store_true")
    ap.add_argument("--ragged", action="store_true", help="random prompt / generation lengths (batch size changes every few steps)")
    ap.add_argument("--max-num-seqs", type=int, default=0, help="batch slots (default: --num-prompts)")
    ap.add_argument("--mp", action="store_true", help="engine core in a subprocess (vLLM default); plugin via entry point")
    ap.add_argument("--tp", type=int, default=1, help="tensor_parallel_size (KV heads sharded over the GPUs of one node)")
    ap.add_argument("--repeats", type=int, default=3, help="timed generate() calls; the median is reported")
    ap.add_argument("--save", default=None)
    ap.add_argument("--compare", nargs=2, default=None)
    args = ap.parse_args()
    if args.compare:
        print(json.dumps(compare(*args.compare)))
        return
    print(json.dumps(run(args)), flush=True)


if __name__ == "__main__":
    main()
