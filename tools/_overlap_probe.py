import sys, torch, time
sys.path.insert(0,'.')
from softtreemax_b200.engine import Engine
from softtreemax_b200.roots import make_roots
eng=Engine("synth_breakout", scratch_envs=65536); dev=eng.device
R,A,d=4096,4,2; L=R*A**d
st,sk=make_roots(eng,R)
def mk(): return dict(stacks=torch.empty((L,4,84,84),dtype=torch.uint8,device=dev),rewards=torch.empty((L,),dtype=torch.float32,device=dev),done=torch.empty((L,),dtype=torch.uint8,device=dev),lives=torch.empty((L,),dtype=torch.int32,device=dev),states=None)
o1,o2=mk(),mk()
eng.cnn_set_weights({k: torch.randn(s)*0.05 for k,s in {"cnn.0.weight":(32,4,8,8),"cnn.0.bias":(32,),"cnn.2.weight":(64,32,4,4),"cnn.2.bias":(64,),"cnn.4.weight":(64,64,3,3),"cnn.4.bias":(64,),"linear.0.weight":(512,3136),"linear.0.bias":(512,),"action_net.weight":(A,512),"action_net.bias":(A,),"value_net.weight":(1,512),"value_net.bias":(1,)}.items()})
eng.bfs(st,sk,d,out=o1); eng.bfs(st,sk,d,out=o2); eng.cnn_forward(o1["stacks"]); torch.cuda.synchronize()
def t(fn,n=4):
    torch.cuda.synchronize(); e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1)/n
print("bfs alone", t(lambda: eng.bfs(st,sk,d,out=o1)))
print("cnn alone", t(lambda: eng.cnn_forward(o2["stacks"])))
sa,sb=torch.cuda.Stream(),torch.cuda.Stream()
def both():
    m=torch.cuda.current_stream(); sa.wait_stream(m); sb.wait_stream(m)
    with torch.cuda.stream(sa): eng.bfs(st,sk,d,out=o1)
    with torch.cuda.stream(sb): eng.cnn_forward(o2["stacks"])
    m.wait_stream(sa); m.wait_stream(sb)
print("both concurrently", t(both))
def both2():
    m=torch.cuda.current_stream(); sa.wait_stream(m); sb.wait_stream(m)
    with torch.cuda.stream(sb): eng.cnn_forward(o2["stacks"])
    with torch.cuda.stream(sa): eng.bfs(st,sk,d,out=o1)
    m.wait_stream(sa); m.wait_stream(sb)
print("both (cnn first)", t(both2))
