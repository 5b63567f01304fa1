for mb in 4 8 16 32 64; do echo "== DNN_DW_MIN_BLOCKS=$mb"; DNN_DW_MIN_BLOCKS=$mb timeout 120 python bench/tc_bwd_probe.py 0 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print({k:v for k,v in d.items() if 'tcgen05' in k and ('N1024' in k or 'N10000' in k or 'N256' in k)})"; done
