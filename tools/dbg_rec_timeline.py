import os, sys, argparse, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ['CRNN_TC_DEBUG']='1'
from distributed_char_rnn_b200.model import Model
E=os.environ.get
ns=argparse.Namespace(model=E('DBG_MODEL','gru'),rnn_size=int(E('DBG_H','512')),num_layers=int(E('DBG_LAYERS','1')),seq_length=int(E('DBG_S','50')),batch_size=int(E('DBG_B','256')),vocab_size=96,grad_clip=5.0)
prec=sys.argv[1] if len(sys.argv)>1 else 'bf16'
m=Model(ns,device='cuda:0',precision=prec,seed=0); m.assign_lr(0.002)
rng=np.random.default_rng(0)
S_=ns.seq_length
x=rng.integers(0,96,(ns.batch_size,S_)).astype(np.int32); y=rng.integers(0,96,(ns.batch_size,S_)).astype(np.int32)
for _ in range(3): m.step(x,y)
torch.cuda.synchronize()
d=m.workspace[24576:24576+8*4*192].view(torch.int64).cpu().numpy().reshape(4,8,24)
names=['prod_start','await_done','tma_issued','mma_issued','epi_at_wait','acc_taken','epi_synced','published','w0_computed','w9_computed','mma_zempty','mma_full_first','mma_full_last','deferred_done','zfull_seen_t0','zfull_seen_t255','zv_in_regs','gates_landed','state_landed','p_empty0','p_issued0','p_empty1','p_issued1','-']
print('precision',prec,'env',{k:v for k,v in os.environ.items() if k.startswith('CRNN_')})
print('stamps are relative to the producer\'s await_done of the same phase (the moment every CTA has published the previous phase)')
order=[0,10,19,20,22,2,11,12,3,4,14,5,16,21,8,6,7,13]
print('cta ph | '+' '.join('%s'%names[i][:12].rjust(12) for i in order)+' | phase_len')
R=int(E('DBG_REF','1'))      # reference stamp (cluster exchange: 11, the moment the MMA warp sees its operand buffer complete)
for c in (0,3):
    for p in range(1,7):
        ref=d[c,p,R]
        print(c,p,'|',' '.join(str(int(d[c,p,i]-ref)).rjust(12) for i in order),'|',int(d[c,p+1,R]-d[c,p,R]))

life=m.workspace[24576+8*4*192:24576+8*4*192+8*16].view(torch.int64).cpu().numpy()
print('CTA 0 life cycle (cycles after entry): setup_done %d  weights_resident %d  second_step_begins %d  mma_loop_end %d  epilogue_loop_end %d  all_warps_done %d  exit %d ; globaltimer entry..exit %.2f us for %d steps'
      % tuple([int(life[i]-life[0]) for i in (1,2,3,4,7,5,6)]+[(life[9]-life[8])/1e3, S_]))
