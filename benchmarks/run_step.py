"""One eager optimizer step of the 124M DDP benchmark config after warm-up: target for
`ncu --metrics gpu__time_duration.sum -s <warm-up launches> -c <launches per step>` (B200_NO_GRAPHS=1 so that every
kernel is its own launch).  Prints the number of native launches per step."""
import os
import sys

os.environ.setdefault("B200_NO_GRAPHS", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from distributed_llm_trainer_b200.models import GPTConfig  # noqa: E402
from distributed_llm_trainer_b200.ops import kernels as K  # noqa: E402
from distributed_llm_trainer_b200.training.config import TrainingConfig  # noqa: E402
from distributed_llm_trainer_b200.training.ddp_trainer import DistributedTrainer  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
tr = DistributedTrainer(GPTConfig.preset("small"), TrainingConfig(batch_size=8, gradient_accumulation_steps=4, max_steps=100,
                                                                  mixed_precision="bf16", seed=1))
ids = torch.randint(0, 50257, (32, 1024)).cuda()
for i in range(steps):
    before = K.launch_count
    tr.train_step({"input_ids": ids}, sync=False)
    torch.cuda.synchronize()
    print("native launches in step", i, K.launch_count - before)
