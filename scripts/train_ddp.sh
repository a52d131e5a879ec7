#!/bin/bash
# DDP launch: ./scripts/train_ddp.sh [NUM_GPUS] [MODEL_SIZE] [extra trainer flags...]
# Same positional interface as the reference's scripts/train_ddp.sh.
set -e
NUM_GPUS=${1:-$(nvidia-smi -L | wc -l)}
MODEL_SIZE=${2:-small}
shift 2 2>/dev/null || shift $# 2>/dev/null
cd "$(dirname "$0")/.."
echo "DDP training | GPUs: $NUM_GPUS | model: $MODEL_SIZE | extra: $*"
export CUDA_VISIBLE_DEVICES=$(seq -s, 0 $((NUM_GPUS-1)))
export OMP_NUM_THREADS=8
python -c "import __graft_entry__ as g; g.build()"
torchrun --standalone --local-addr 127.0.0.1 --nproc_per_node="$NUM_GPUS" src/training/ddp_trainer.py \
    --model_size "$MODEL_SIZE" --batch_size 8 --max_steps 1000 --mixed_precision bf16 "$@"
echo "Training complete!"
