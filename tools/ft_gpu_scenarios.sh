#!/bin/bash
# Fault-tolerance scenarios on real GPUs, through train.py and a fake sbatch (records its argv).
#   tools/ft_gpu_scenarios.sh peer   2-GPU run, rank 1 raises on its own at step 6 -> both ranks save step 6, resume
#   tools/ft_gpu_scenarios.sh fsdpN  N-GPU (default 8) fully-sharded runs of the FULL 8.05 B model (BASELINE.json config 4):
#                                    injected exception at step 30 -> every rank saves step 30, no resubmit;
#                                    SIGTERM to every rank -> "Job cancelled", nothing written
#   tools/ft_gpu_scenarios.sh logs   1-GPU runs of the FULL 8.05 B model reproducing the reference's three evidence logs
#                                    (timeout -> save + resubmit; resume + injected error -> save; scancel -> no save)
set -u
cd "$(dirname "$0")/.."
W=/tmp/ftl_ft; rm -rf $W; mkdir -p $W/bin gpurun_out/logs
printf '#!/bin/bash\necho "$@" >> %s/sbatch_calls.txt\necho "Submitted batch job 900002"\n' $W > $W/bin/sbatch; chmod +x $W/bin/sbatch
export PATH=$W/bin:$PATH WORKDIR=$W PYTHONUNBUFFERED=1
TINY="--tokenizer-name-or-path offline --vocab-size 8192 --dim 1024 --n-layers 3 --n-heads 8 --n-kv-heads 4 --multiple-of 256 --sequence-length 256 --batch-size 1 --dataset $W/d.parquet --synthetic-docs 64 --learning-rate 1e-3 --lr-warmup-steps 5 --logging-frequency 1"
run2() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 train.py "${@:2}"; }
if [ "${1#fsdp}" != "$1" ]; then
  NP=${1#fsdp}; NP=${NP:-8}
  runN() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port $1 train.py "${@:2}"; }
  FULLF="--no-consolidate --parallel fsdp --tokenizer-name-or-path offline --sequence-length 2048 --batch-size 1 --learning-rate 5e-5 --lr-warmup-steps 100 --dataset $W/full.parquet --synthetic-docs 512 --logging-frequency 5"
  echo "== job 910001: FSDP x$NP, --raise-error --error-step 30"
  SLURM_JOB_ID=910001 runN 29561 $FULLF --training-steps 200 --raise-error --error-step 30 > gpurun_out/logs/fsdp${NP}_error_step.out 2>&1; echo "rc=$?"
  grep -E "EXIT HANDLER|Training step: (0|25|30) " gpurun_out/logs/fsdp${NP}_error_step.out | cut -c1-200
  ls -la $W/checkpoints/ 2>/dev/null | head; du -sh $W/checkpoints/* 2>/dev/null | head -3
  [ -f $W/sbatch_calls.txt ] && echo "sbatch calls: $(cat $W/sbatch_calls.txt)" || echo "sbatch calls: none"
  rm -rf $W/checkpoints
  echo "== job 910002: FSDP x$NP, SIGTERM (scancel) after ~15 steps"
  SLURM_JOB_ID=910002 runN 29562 $FULLF --training-steps 100000 > gpurun_out/logs/fsdp${NP}_sigterm.out 2>&1 &
  PID=$!
  for i in $(seq 1 600); do grep -q "Training step: 15 " gpurun_out/logs/fsdp${NP}_sigterm.out 2>/dev/null && break; sleep 0.5; kill -0 $PID 2>/dev/null || break; done
  for p in $(pgrep -P $PID); do kill -TERM $p 2>/dev/null; done     # the job step's tasks, as srun / scancel would
  wait $PID; echo "rc=$?"
  grep -E "EXIT HANDLER|Training step: 15 " gpurun_out/logs/fsdp${NP}_sigterm.out | cut -c1-200
  ls $W/checkpoints 2>/dev/null | head -3; echo "checkpoint dir entries: $(ls $W/checkpoints 2>/dev/null | wc -l)"
elif [ "$1" = "peer" ]; then
  for mode in dp fsdp; do
    echo "== $mode: rank-local failure on rank 1 at step 6"
    SLURM_JOB_ID=81$mode FTL_INJECT_RANK_ERROR=1:6 run2 29551 $TINY --training-steps 14 --parallel $mode > gpurun_out/logs/peer_$mode.log 2>&1
    grep -E "injected|EXIT HANDLER|rewinding" gpurun_out/logs/peer_$mode.log | cut -c1-220
    echo "== $mode: resume"
    SLURM_JOB_ID=82$mode run2 29552 $TINY --training-steps 14 --parallel $mode --checkpoint-id 81$mode > gpurun_out/logs/peer_${mode}_resume.log 2>&1
    grep -E "Resuming|Training completed|Training step: 13" gpurun_out/logs/peer_${mode}_resume.log | cut -c1-200
  done
else
  # the box has ~80 GB of free disk: one 48 GB copy at a time, so shards only (--no-consolidate) and the first job's
  # checkpoint is removed once the second job has loaded it
  FULL="--no-consolidate --tokenizer-name-or-path offline --sequence-length 2048 --batch-size 1 --learning-rate 5e-5 --lr-warmup-steps 100 --dataset $W/full.parquet --synthetic-docs 128 --raise-error --error-step 60"
  echo "== job 900001: fresh start, SIGUSR1 after ~25 steps"
  SLURM_JOB_ID=900001 python train.py $FULL --training-steps 1000 > gpurun_out/logs/output_900001.out 2>&1 &
  PID=$!
  while ! grep -q "Training step: 25 " gpurun_out/logs/output_900001.out 2>/dev/null; do sleep 0.5; kill -0 $PID 2>/dev/null || break; done
  kill -USR1 $PID; wait $PID; echo "rc=$?"
  echo "== job 900002: resume from 900001, injected exception at step 60"
  SLURM_JOB_ID=900002 python train.py $FULL --training-steps 1000 900001 > gpurun_out/logs/output_900002.out 2>&1 &
  PID=$!
  while ! grep -q "Resuming training" gpurun_out/logs/output_900002.out 2>/dev/null; do sleep 0.5; kill -0 $PID 2>/dev/null || break; done
  rm -rf $W/checkpoints/checkpoint_900001.shards; wait $PID; echo "rc=$?"
  echo "== job 900003: fresh start, scancel (SIGTERM) after ~15 steps"
  SLURM_JOB_ID=900003 python train.py $FULL --training-steps 1000 > gpurun_out/logs/output_900003.out 2>&1 &
  PID=$!
  while ! grep -q "Training step: 15 " gpurun_out/logs/output_900003.out 2>/dev/null; do sleep 0.5; kill -0 $PID 2>/dev/null || break; done
  kill -TERM $PID; wait $PID; echo "rc=$?"
  cat $W/sbatch_calls.txt; ls -la $W/checkpoints/ | head
  for f in 900001 900002 900003; do echo "--- $f"; grep -E "EXIT HANDLER|Resuming|Checkpoint read" gpurun_out/logs/output_$f.out | cut -c1-200; done
fi
