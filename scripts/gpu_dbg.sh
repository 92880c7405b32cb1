mkdir -p gpurun_out
for v in "" "8192,1024,1,1,1024,1,4" "8192,512,1,1,512,1,4" "8192,512,1,1,1024,1,3" "8192,1024,0,0,1024,2,3"; do
  if [ -n "$v" ]; then export DCNN_WGRAD_DBG="$v"; fi
  timeout 120 python scripts/dbg_wgrad.py 2>&1 | grep -E "^\[|DBG|Error|error"
  echo "-----------------"
done > gpurun_out/dbg_wgrad.log 2>&1
tail -60 gpurun_out/dbg_wgrad.log
