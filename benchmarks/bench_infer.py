"""BASELINE config 5: GPT-2 "124M" greedy generation (50 new tokens) from a checkpoint WRITTEN BY THE NEW TRAINER.

1. `src/training/ddp_trainer.py --model_size small --dataset dummy --max_steps N` (the reference's CLI path) writes
   checkpoints/final.pt on the GPU;
2. `src/eval/infer.py --checkpoint ... --greedy --max_new_tokens 50 --prompt_ids ...` generates from it (this repo's
   engine: every token re-runs the fused forward, last-position LM head only -- no KV cache, as the reference);
3. the same file is loaded into the UNMODIFIED reference `GPT` (baseline/_ref) and generated from with its own
   `generate()` restricted to arg-max (top_k=1), on the GPU, and the token ids are compared;
4. both are timed (CUDA events around generate, after one warm-up generate).

    python benchmarks/bench_infer.py [--steps 20]
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed_generate(fn, torch):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    out = fn()
    b.record()
    torch.cuda.synchronize()
    return out, a.elapsed_time(b)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--new_tokens", type=int, default=50)
    args = ap.parse_args()
    import torch

    tmp = tempfile.mkdtemp(prefix="b200_infer_")
    env = dict(os.environ, PYTHONPATH=ROOT, MASTER_ADDR="127.0.0.1")
    t0 = time.time()
    r = subprocess.run([sys.executable, os.path.join(ROOT, "src", "training", "ddp_trainer.py"), "--model_size", "small",
                        "--dataset", "dummy", "--max_steps", str(args.steps), "--batch_size", "8", "--checkpoint_dir", tmp,
                        "--seed", "1234"], cwd=ROOT, env=env, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    ckpt = os.path.join(tmp, "final.pt")
    assert os.path.exists(ckpt), os.listdir(tmp)
    train_s = time.time() - t0
    prompt = [464, 2068, 7586, 21831, 11687, 625, 262, 16931, 3290, 13]

    from distributed_llm_trainer_b200.eval.infer import load_model

    dev = torch.device("cuda")
    model = load_model(ckpt, None, dev)
    ids = torch.tensor([prompt], device=dev)
    ours, ms_ours = timed_generate(lambda: model.generate(ids, max_new_tokens=args.new_tokens, greedy=True), torch)
    res = {"checkpoint_bytes": os.path.getsize(ckpt), "train_steps": args.steps, "train_wall_s": round(train_s, 1),
           "new_tokens": args.new_tokens, "b200": {"ms": ms_ours, "tokens_per_s": args.new_tokens / ms_ours * 1e3,
                                                   "token_ids": ours[0].tolist()}}
    del model
    torch.cuda.empty_cache()
    # the unmodified reference model loads the same file
    ref_root = os.path.join(ROOT, "baseline", "_ref")
    if os.path.isdir(os.path.join(ref_root, "models")):
        for m in [k for k in sys.modules if k.split(".")[0] in ("models", "training", "data", "eval")]:
            del sys.modules[m]
        sys.path.insert(0, ref_root)
        from models.config import GPTConfig as RefCfg
        from models.gpt import GPT as RefGPT

        ck = torch.load(ckpt, map_location="cpu", weights_only=False)
        rmodel = RefGPT(RefCfg.gpt2_small())
        rmodel.load_state_dict(ck["model"])            # strict: identical state-dict layout
        rmodel = rmodel.to(dev).eval()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            ref, ms_ref = timed_generate(lambda: rmodel.generate(ids, max_new_tokens=args.new_tokens, top_k=1), torch)
        agree = sum(int(a == b) for a, b in zip(ours[0].tolist(), ref[0].tolist())) / ours.shape[1]
        res["reference"] = {"ms": ms_ref, "tokens_per_s": args.new_tokens / ms_ref * 1e3, "token_ids": ref[0].tolist(),
                            "loaded_strict": True, "note": "reference GPT.generate(top_k=1) under bf16 autocast"}
        res["token_agreement"] = agree
        res["speedup"] = ms_ref / ms_ours
    print(json.dumps(res))


if __name__ == "__main__":
    main()
