#!/bin/bash
# round 2, GPU call 16: ncu --set full of the packed 16-bit pooling kernel (c1 bf16) next to the index kernel (c1 fp32)
o=gpurun_out
python tools/bench_ops.py --config c1 --dtype bf16 --iters 2 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"roi_pool_fwd_rmqp" -s 2 -c 1 -o $o/r2p_pack16 python tools/bench_ops.py --config c1 --dtype bf16 --iters 2 > $o/r2p_ncu_pack.log 2>&1
python tools/bench_ops.py --config c1 --iters 2 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"roi_pool_fwd_rmq_kernel" -s 2 -c 1 -o $o/r2p_index_c1 python tools/bench_ops.py --config c1 --iters 2 > $o/r2p_ncu_index.log 2>&1
ls -la $o | grep r2p
