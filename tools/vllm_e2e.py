#!/usr/bin/env python3
"""End-to-end decode through vLLM's own engine, with and without the b200c plugin.

    python tools/vllm_e2e.py --tp 2 --model-shape llama3-8b --batch 64 --decode 128 --plugin on
Random-init bf16 weights of the named shape (`load_format=dummy`, fixed seed), random prompt token
ids (seeded), greedy sampling.  Prints one JSON line: decode tokens/s, prefill time, and a digest
of the generated token ids so two runs (plugin on / off) can be compared for identity.
The reference launcher (launch.py) is not used here: it does not import against vLLM 0.22
(SURVEY.md appendix A); vLLM's own multiprocess executor starts the same WorkerWrapperBase
workers, and the plugin is loaded in each of them exactly as it would be under launch.py.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SHAPES = {
    "llama3-8b": dict(hidden_size=4096, intermediate_size=14336, num_hidden_layers=32,
                      num_attention_heads=32, num_key_value_heads=8, vocab_size=128256,
                      rms_norm_eps=1e-5, rope_theta=500000.0, max_position_embeddings=8192),
    "llama3-70b": dict(hidden_size=8192, intermediate_size=28672, num_hidden_layers=80,
                       num_attention_heads=64, num_key_value_heads=8, vocab_size=128256,
                       rms_norm_eps=1e-5, rope_theta=500000.0, max_position_embeddings=8192),
    "tiny": dict(hidden_size=4096, intermediate_size=14336, num_hidden_layers=4,
                 num_attention_heads=32, num_key_value_heads=8, vocab_size=128256,
                 rms_norm_eps=1e-5, rope_theta=500000.0, max_position_embeddings=8192),
}


def write_model_dir(shape: dict, layers: int | None) -> str:
    cfg = dict(architectures=["LlamaForCausalLM"], model_type="llama", hidden_act="silu",
               torch_dtype="bfloat16", tie_word_embeddings=False, bos_token_id=128000,
               eos_token_id=128001, attention_bias=False, mlp_bias=False, **shape)
    if layers:
        cfg["num_hidden_layers"] = layers
    d = tempfile.mkdtemp(prefix="b200c_model_")
    json.dump(cfg, open(os.path.join(d, "config.json"), "w"))
    return d


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tp", type=int, default=2)
    ap.add_argument("--pp", type=int, default=1)
    ap.add_argument("--model-shape", default="llama3-8b", choices=sorted(SHAPES))
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--batch", type=int, nargs="+", default=[64])
    ap.add_argument("--prompt-len", type=int, default=1024)
    ap.add_argument("--decode", type=int, default=128)
    ap.add_argument("--plugin", default="on", choices=["on", "off"])
    ap.add_argument("--fuse-layers", default="1")
    ap.add_argument("--eager", action="store_true")
    ap.add_argument("--stock-fusion", action="store_true",
                    help="plugin off: leave vLLM's own FlashInfer allreduce+rms fusion pass on")
    ap.add_argument("--out", default="")
    args = ap.parse_args()

    os.environ["PYTHONPATH"] = ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")
    sys.path.insert(0, ROOT)
    os.chdir(ROOT)
    os.environ["B200C_DISABLE"] = "0" if args.plugin == "on" else "1"
    os.environ["B200C_FUSE_LAYERS"] = args.fuse_layers
    os.environ.setdefault("VLLM_WORKER_MULTIPROC_METHOD", "spawn")
    os.environ.setdefault("VLLM_LOGGING_LEVEL", "INFO")
    fi_aot = os.path.join(ROOT, "baseline", "_fi", "aot")
    if os.path.isdir(fi_aot):
        # the stock comparator's FlashInfer module, cross-compiled ahead of time (bench infrastructure)
        shim = os.path.join(ROOT, "tools", "fi_jit_cache_shim")
        os.environ["PYTHONPATH"] = shim + os.pathsep + os.environ["PYTHONPATH"]
        os.environ["FLASHINFER_DISABLE_VERSION_CHECK"] = "1"
        sys.path.insert(0, shim)

    import random
    from vllm import LLM, SamplingParams

    model_dir = write_model_dir(SHAPES[args.model_shape], args.layers)
    comp = {}
    if args.plugin == "on" or not args.stock_fusion:
        comp = {"pass_config": {"fuse_allreduce_rms": False}}
    t0 = time.time()
    llm = LLM(model=model_dir, tensor_parallel_size=args.tp, pipeline_parallel_size=args.pp,
              load_format="dummy", skip_tokenizer_init=True, dtype="bfloat16", seed=0,
              max_model_len=args.prompt_len + args.decode + 16, gpu_memory_utilization=0.85,
              enforce_eager=args.eager, max_num_seqs=max(max(args.batch), 8),
              max_num_batched_tokens=max(8192, args.prompt_len),
              distributed_executor_backend="mp" if args.tp * args.pp > 1 else None,
              compilation_config=comp or None, enable_prefix_caching=False)
    t_init = time.time() - t0
    vocab = SHAPES[args.model_shape]["vocab_size"]
    all_tokens = {}
    for batch in args.batch:
        rng = random.Random(1234)
        prompts = [{"prompt_token_ids": [rng.randrange(1000, vocab - 1000) for _ in range(args.prompt_len)]}
                   for _ in range(batch)]

        def run(max_tokens):
            sp = SamplingParams(temperature=0.0, max_tokens=max_tokens, ignore_eos=True, detokenize=False)
            t = time.time()
            outs = llm.generate(prompts, sp, use_tqdm=False)
            return time.time() - t, [list(o.outputs[0].token_ids) for o in outs]

        run(4)  # warm-up (graph replay paths, allocator)
        t_prefill, _ = run(1)
        t_full, toks = run(args.decode)
        t_decode = max(1e-9, t_full - t_prefill)
        digest = hashlib.sha256(json.dumps(toks).encode()).hexdigest()[:16]
        res = {"tp": args.tp, "pp": args.pp, "shape": args.model_shape, "layers": args.layers or
               SHAPES[args.model_shape]["num_hidden_layers"], "batch": batch,
               "prompt_len": args.prompt_len, "decode": args.decode, "plugin": args.plugin,
               "stock_fusion": args.stock_fusion, "fuse_layers": args.fuse_layers, "eager": args.eager,
               "decode_tokens_per_s": round(batch * (args.decode - 1) / t_decode, 1),
               "prefill_s": round(t_prefill, 4), "full_s": round(t_full, 4),
               "init_s": round(t_init, 1), "digest": digest}
        print("E2E_RESULT " + json.dumps(res), flush=True)
        all_tokens[str(batch)] = {"result": res, "tokens": toks}
    if args.out:
        json.dump(all_tokens, open(args.out, "w"))
    del llm


if __name__ == "__main__":
    main()
