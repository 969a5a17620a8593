python scripts/time_longform.py 2>/dev/null
cat > /tmp/lf1.py <<'PY'
import sys, os, torch
sys.path.insert(0, os.getcwd())
import paper_accurate_fast_cheap_b200 as pk
T,H,N=30000,8,64; C=H*N; dev="cuda:0"
g=torch.Generator(device=dev).manual_seed(2)
def mk(B):
    r,k,v=(torch.randn(B,T,C,device=dev,generator=g).bfloat16() for _ in range(3))
    w=(-6+5*torch.rand(B,T,C,device=dev,generator=g)).bfloat16(); u=(0.3*torch.randn(H,N,device=dev,generator=g)).bfloat16()
    return r,k,v,w,u
f,b=mk(1),mk(1)
for _ in range(3): pk.bi_forward_raw(f,b)
torch.cuda.synchronize()
PY
ncu --metrics gpu__time_duration.sum --clock-control none --csv python /tmp/lf1.py 2>/dev/null | grep -i "wkv6" | awk -F'","' '{print $5, $NF}' | tail -6
