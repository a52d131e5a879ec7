#!/bin/bash
# FSDP launch: ./scripts/train_fsdp.sh [NUM_GPUS] [MODEL_SIZE] [extra flags, e.g. --cpu_offload]
# Same positional interface as the reference's scripts/train_fsdp.sh.
set -e
NUM_GPUS=${1:-$(nvidia-smi -L | wc -l)}
MODEL_SIZE=${2:-medium}
shift 2 2>/dev/null || shift $# 2>/dev/null
cd "$(dirname "$0")/.."
echo "FSDP training | GPUs: $NUM_GPUS | model: $MODEL_SIZE | extra: $*"
export CUDA_VISIBLE_DEVICES=$(seq -s, 0 $((NUM_GPUS-1)))
export OMP_NUM_THREADS=8
export NCCL_DEBUG=WARN
python -c "import __graft_entry__ as g; g.build()"
torchrun --standalone --local-addr 127.0.0.1 --nproc_per_node="$NUM_GPUS" src/training/fsdp_trainer.py \
    --model_size "$MODEL_SIZE" --batch_size 4 --max_steps 1000 --sharding FULL_SHARD "$@"
echo "Training complete!"
