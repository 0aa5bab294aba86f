#!/bin/bash
# usage: scripts/retry_gpu.sh <script-or-bash-args> <outfile> : runs `bash <script>` on the GPU box, retrying while the pool is busy
for i in $(seq 1 12); do
  /usr/local/graft/bin/gpurun --timeout 2400 -- "bash $1" > "$2" 2>&1
  if grep -q "status=ok" "$2"; then exit 0; fi
  if ! grep -q "status=transient\|exit code 3\|no box" "$2"; then exit 1; fi
  sleep 150
done
exit 2
