import os, sys, argparse, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ['CRNN_TC_DEBUG']='1'
from distributed_char_rnn_b200.model import Model
ns=argparse.Namespace(model='gru',rnn_size=512,num_layers=1,seq_length=50,batch_size=256,vocab_size=96,grad_clip=5.0)
prec=sys.argv[1] if len(sys.argv)>1 else 'bf16'
m=Model(ns,device='cuda:0',precision=prec,seed=0); m.assign_lr(0.002)
rng=np.random.default_rng(0)
x=rng.integers(0,96,(256,50)).astype(np.int32); y=rng.integers(0,96,(256,50)).astype(np.int32)
for _ in range(3): m.step(x,y)
torch.cuda.synchronize()
d=m.workspace[24576:24576+8*16*60].view(torch.int64).cpu().numpy().reshape(60,16)
# phases 20..39
base=d[20,0]
names=['prod_start','await_done','tma_issued','mma_done','epi_loaded','acc_taken','epi_synced','published','w0_computed','w3_computed','mma_zempty','mma_full0','mma_full1']
print(names)
print('precision',prec,'env',{k:v for k,v in os.environ.items() if k.startswith('CRNN_')})
for p in range(20,28):
    print(p, [int(v-d[p,0]) for v in d[p][:13]], 'phase_len', int(d[p+1,0]-d[p,0]))
