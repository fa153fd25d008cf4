#!/bin/bash
# Launch one rank of a multi-rank bench.py run; rank 0 alone runs under ncu (never wrap a whole
# multi-rank command: ncu serialises and may replay kernels).  One pass per kernel only -- NVLink
# byte counters -- because a collective kernel cannot be replayed: its peers would not resend.
#   python -m torch.distributed.run --no-python --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
#       --master-port 29514 bash tools/ncu_rank0.sh gpurun_out/nvlink_rank0.csv --gpus 2 --steps 2 --warmup 3 --no-extras
out="$1"; shift
if [ "${LOCAL_RANK:-0}" = "0" ]; then
  exec ncu --metrics nvlrx__bytes.sum,nvltx__bytes.sum --clock-control none --cache-control none \
       -k regex:"ar_norm|ar_kernel|all_gather|nvls" -c 60 --csv --log-file "$out" python bench.py "$@"
else
  exec python bench.py "$@"
fi
