#!/bin/bash
# Slurm launcher with the reference's contract (reference: train.sh:1-30):
#   sbatch train.sh            fresh run
#   sbatch train.sh <JOBID>    resume from checkpoint_<JOBID>.ckpt
# Slurm sends USR1 120 s before the time limit (--signal=USR1@120); the trainer then
# checkpoints and calls `sbatch $WORKDIR/train.sh $SLURM_JOB_ID` itself (--no-requeue).
# Differences from the reference: one task per GPU on one 8xB200 node, launched by srun with
# RANK/LOCAL_RANK/WORLD_SIZE derived from Slurm's variables; GPUS can be overridden.
#SBATCH --job-name=ftl_b200
#SBATCH --nodes=1
#SBATCH --ntasks-per-node=8
#SBATCH --gpus-per-node=8
#SBATCH --time=00:06:00
#SBATCH --signal=USR1@120
#SBATCH --no-requeue

TRAINING_CMD=" --sequence-length 2048 \
               --batch-size 1 \
               --learning-rate 5e-5 \
               --lr-warmup-steps 100 \
               --training-steps 1000 \
               --raise-error \
               --error-step 600 \
               ${FTL_EXTRA_ARGS:-}"

if [ -n "$1" ]; then
    TRAINING_CMD="$TRAINING_CMD --checkpoint-id $1"
fi
export WORKDIR="${WORKDIR:-$(cd "$(dirname "$0")" && pwd)}"
export MASTER_ADDR="${MASTER_ADDR:-127.0.0.1}"
export MASTER_PORT="${MASTER_PORT:-29511}"

if command -v srun >/dev/null 2>&1 && [ -n "$SLURM_JOB_ID" ]; then
    # srun forwards USR1/TERM to every task
    exec srun --unbuffered bash -c 'export RANK=$SLURM_PROCID LOCAL_RANK=$SLURM_LOCALID WORLD_SIZE=$SLURM_NTASKS; \
        exec python '"$WORKDIR"'/train.py '"$TRAINING_CMD"
else
    GPUS="${GPUS:-1}"
    if [ "$GPUS" -gt 1 ]; then
        exec python -m torch.distributed.run --nnodes=1 --nproc-per-node "$GPUS" \
            --master-addr "$MASTER_ADDR" --master-port "$MASTER_PORT" "$WORKDIR/train.py" $TRAINING_CMD
    fi
    exec python "$WORKDIR/train.py" $TRAINING_CMD
fi
