#!/usr/bin/env bash
# generation-3 bring-up: quick check on small shapes, full parity suite with FA_SM100_GEN=3, sweep, trace
set -u
mkdir -p gpurun_out
export FA_SM100_GEN=3
timeout 180 python - <<'PY' 2>&1 | tee gpurun_out/quick.log
import sys, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from common import errors
from oracle.eager_attention import eager_attention
from flash_attention_cute_b200 import _lib
from flash_attention_cute_b200.load_cpp_extention import load_extension
ext = load_extension(); dev = torch.device('cuda:0')
for (B,Hq,Hkv,Sq,Sk,D,causal) in [(1,1,1,128,128,128,False),(1,1,1,128,256,128,False),(1,1,1,128,512,128,False),(1,2,2,512,512,128,True),(2,8,2,1024,1024,128,True),(1,4,4,1000,1000,64,True),(4,16,16,2048,2048,128,False),(1,14,2,1,512,64,True),(1,4,4,300,200,64,True)]:
    g = torch.Generator(device=dev).manual_seed(1)
    q = torch.randn(B,Hq,Sq,D,generator=g,device=dev).bfloat16(); k = torch.randn(B,Hkv,Sk,D,generator=g,device=dev).bfloat16(); v = torch.randn(B,Hkv,Sk,D,generator=g,device=dev).bfloat16()
    try:
        o = ext.flash_attention_fwd(q,k,v,D**-0.5,causal,0); _lib.check_device(0)
        ref = eager_attention(q,k,v,None,causal,out_dtype=torch.float32)
        print((B,Hq,Hkv,Sq,Sk,D,causal), 'max_abs %.3e rel_l2 %.3e' % errors(o, ref), 'nan', int(torch.isnan(o.float()).sum()), flush=True)
    except Exception as e:
        print((B,Hq,Hkv,Sq,Sk,D,causal), 'ERROR', repr(e)[:300], flush=True); break
PY
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_gen3.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu_gen3.log
unset FA_SM100_GEN
timeout 900 python tools/tune.py --iters 10 --shapes c3b2,c3,c2,c3nc,c4s4k --halves 1,3 --emu 0,1,2 --split 128 2>&1 | tee gpurun_out/tune.log
timeout 120 python tools/trace_view.py c3nc --halves 3 --emu 0 --dump gpurun_out/trace_c3nc_gen3.json 2>&1 | tee gpurun_out/trace.log
