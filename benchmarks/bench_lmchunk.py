"""End-to-end step time vs LM-head chunk size (bytes of the bf16 logits chunk)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_trainer_b200.models import GPT, GPTConfig  # noqa: E402
from distributed_llm_trainer_b200.ops import engine as E  # noqa: E402

for mb in (192, 288, 420, 900):
    E._LM_CHUNK_BYTES = mb << 20
    torch.manual_seed(0)
    model = GPT(GPTConfig.gpt2_small()).cuda()
    eng = model.engine()
    ids = torch.randint(0, 50257, (8, 1024), device="cuda")
    for _ in range(4):
        eng.forward_backward(ids, ids, 0.25)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(12):
        eng.forward_backward(ids, ids, 0.25)
    b.record()
    torch.cuda.synchronize()
    ws = next(iter(eng._ws.values()))
    print(f"chunk {mb} MB -> {ws.lm_rows} rows: {a.elapsed_time(b)/12:.3f} ms per micro-step")
    del model, eng
    torch.cuda.empty_cache()
