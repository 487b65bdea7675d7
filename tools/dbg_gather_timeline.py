# Per-phase clock64 stamps of the forward recurrence kernel in direct-exchange mode (CRNN_TC_DEBUG=1, one layer, one launch).
import os, sys, argparse, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ['CRNN_TC_DEBUG'] = '1'
from distributed_char_rnn_b200.model import Model
model = os.environ.get('DBG_MODEL', 'gru'); H = int(os.environ.get('DBG_H', '512')); B = int(os.environ.get('DBG_B', '256'))
ns = argparse.Namespace(model=model, rnn_size=H, num_layers=int(os.environ.get('DBG_LAYERS', '1')), seq_length=int(os.environ.get('DBG_S', '50')), batch_size=B, vocab_size=96, grad_clip=5.0)
prec = sys.argv[1] if len(sys.argv) > 1 else 'bf16'
m = Model(ns, device='cuda:0', precision=prec, seed=0); m.assign_lr(0.002)
rng = np.random.default_rng(0)
S_ = ns.seq_length
x = rng.integers(0, 96, (B, S_)).astype(np.int32); y = rng.integers(0, 96, (B, S_)).astype(np.int32)
for _ in range(3): m.step(x, y)
torch.cuda.synchronize()
d = m.workspace[24576:24576 + 8 * 4 * 192].view(torch.int64).cpu().numpy().reshape(4, 8, 24)
print('precision', prec, 'env', {k: v for k, v in os.environ.items() if k.startswith('CRNN_')})
cols = [('probe_done', 1), ('tma_issued', 2), ('v_start', 19), ('v_landed0', 20), ('v_valid0', 21), ('v_done', 22), ('mma_zempty', 10), ('mma_full0', 11), ('mma_fullN', 12), ('mma_issued', 3),
        ('epi_wait', 4), ('zfull_t0', 14), ('acc_taken', 5), ('zv_regs', 16), ('gates_ld', 17), ('state_ld', 18), ('computed', 8), ('bf16_out', 6), ('deferred', 13)]
print('stamps relative to the producer warp reaching the phase (prod_start); retries = re-reads by thread 0')
print('cta ph | ' + ' '.join(n.rjust(10) for n, _ in cols) + ' |  retries  phase_len')
for c in (0, 3):
    for p in range(1, 7):
        ref = d[c, p, 0]
        print(c, p, '|', ' '.join(str(int(d[c, p, i] - ref)).rjust(10) for _, i in cols), '|', str(int(d[c, p, 23])).rjust(8), int(d[c, p + 1, 0] - d[c, p, 0]))
