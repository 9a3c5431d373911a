"""Time K4 (stm_cnn_forward) alone on 32 768 random observation stacks.

STM_CNN_DEBUG (a bit mask read by stm_cnn_forward, timing experiments only — results are wrong when set) switches
stages of the kernel off to see what the rest costs: 1 producers skip the global loads, 2 the MMA thread skips the
MMAs, 4 the epilogue skips the stores, 8 run conv1 only, 16 conv1 producers skip convert + store, 32 the epilogue
skips the TMEM loads, 64 no proxy fence, 128 plain mbarrier arrives instead of tcgen05.commit.

    STM_CNN_DEBUG=8 python tools/cnn_probe.py
"""
import sys, torch, os
sys.path.insert(0,'.')
from softtreemax_b200.engine import Engine
eng=Engine("synth_breakout"); dev=eng.device
import os as _os
A=4; n=int(_os.environ.get("N_IMAGES", "32768"))
eng.cnn_set_weights({k: torch.randn(s)*0.05 for k,s in {"cnn.0.weight":(32,4,8,8),"cnn.0.bias":(32,),"cnn.2.weight":(64,32,4,4),"cnn.2.bias":(64,),"cnn.4.weight":(64,64,3,3),"cnn.4.bias":(64,),"linear.0.weight":(512,3136),"linear.0.bias":(512,),"action_net.weight":(A,512),"action_net.bias":(A,),"value_net.weight":(1,512),"value_net.bias":(1,)}.items()})
obs=torch.randint(0,256,(n,4,84,84),dtype=torch.uint8,device=dev)
for _ in range(2): eng.cnn_forward(obs)
torch.cuda.synchronize()
e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True); e0.record()
for _ in range(5): eng.cnn_forward(obs)
e1.record(); torch.cuda.synchronize(); print("STM_CNN_DEBUG=%s"%os.environ.get("STM_CNN_DEBUG","0"), "ms per %d images:" % n, e0.elapsed_time(e1)/5)
