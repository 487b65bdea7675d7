import os, sys, argparse, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ['CRNN_TC_CLUSTER'] = sys.argv[1] if len(sys.argv) > 1 else '16'
os.environ['CRNN_TC_HC'] = '32'
from distributed_char_rnn_b200.model import Model
ns = argparse.Namespace(model='gru', rnn_size=512, num_layers=2, seq_length=100, batch_size=256, vocab_size=96, grad_clip=5.0)
m = Model(ns, device='cuda:0', precision='bf16', seed=0); m.assign_lr(0.002)
rng = np.random.default_rng(0)
x = rng.integers(0, 96, (256, 100)).astype(np.int32); y = rng.integers(0, 96, (256, 100)).astype(np.int32)
try:
    for _ in range(3): print(m.step(x, y)[0])
    torch.cuda.synchronize()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(10): m.step_async(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda())
    t1.record(); torch.cuda.synchronize()
    print('ms/step', t0.elapsed_time(t1) / 10)
except Exception as e:
    print('FAILED:', e)
