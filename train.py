#!/usr/bin/env python
"""Entry point with the reference's command line (reference: train.py:131-134).

    python train.py --sequence-length 2048 --batch-size 1 --learning-rate 5e-5 \
        --lr-warmup-steps 100 --training-steps 1000 [--raise-error --error-step 600] \
        [--checkpoint-id <JOBID> | --checkpoint_id=<JOBID> | <JOBID>]

Multi-GPU: `torchrun --nproc-per-node N train.py ... --parallel {dp,fsdp}` (or srun with
RANK/LOCAL_RANK/WORLD_SIZE exported).
"""
from fault_tolerant_llm_training_b200.cli import get_args
from fault_tolerant_llm_training_b200.log import init_logger
from fault_tolerant_llm_training_b200.train_loop import train

if __name__ == "__main__":
    init_logger()
    train(get_args())
