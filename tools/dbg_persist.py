import os, sys, argparse, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_char_rnn_b200.model import Model
ns = argparse.Namespace(model='lstm', rnn_size=32, num_layers=2, seq_length=6, batch_size=5, vocab_size=11, grad_clip=5.0)
m = Model(ns, device='cuda:0', precision='bf16', seed=0)
rng = np.random.default_rng(0)
x = rng.integers(0, 11, (5, 6)).astype(np.int32); y = rng.integers(0, 11, (5, 6)).astype(np.int32)
mode = sys.argv[1] if len(sys.argv) > 1 else 'fb'
if mode == 'fwd':
    m.forward(x, m.zero_state()); torch.cuda.synchronize(); print('forward ok')
else:
    loss, _ = m.gradients(x, y, m.zero_state()); print('fb ok', loss)
