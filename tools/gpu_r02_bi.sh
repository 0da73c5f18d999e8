#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 tools/tp_parity_diag.py"
WF_TP_PARITY_BATCHES=1,16,16,16,16,16,16,16,16,16,16,16,16,16,16,16,16,16,16,16,16,20,20 timeout 200 $T 2>&1 | grep DIAG | python -c "
import sys,ast
for line in sys.stdin:
    i=line.index('['); d=ast.literal_eval(line[i:])
    print('default tree (TP engines: stream off; TP = 1 reference: stream on): cases', len(d), 'mismatches', sum(1 for c in d if not c['bit_identical']), [c['batch'] for c in d if not c['bit_identical']])
" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
