run() { JRL_SLAB_ORDER=$1 JRL_SLAB_GRAPH=$2 timeout 60 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $3 tools/trace_slab.py --no-trace --workload $4 --rows-per-rank $5 --width $6 2>&1 | grep '^{' ; }
run exchange_first 0 29601 c4 1024 8192
run interior_first 0 29602 c4 1024 8192
run exchange_first 1 29603 c4 1024 8192
run interior_first 1 29604 c4 1024 8192
run exchange_first 0 29605 bgk 4096 32768
run interior_first 0 29606 bgk 4096 32768
