# quick device-resident timing of one config in given precision(s): python quick_bench.py c3 bf16x3,bf16 [steps]
import os, sys, argparse, time, numpy as np, torch
sys.path.insert(0, os.getcwd())
import bench as BN
from distributed_char_rnn_b200.model import Model
cname = sys.argv[1]; precs = sys.argv[2].split(','); K = int(sys.argv[3]) if len(sys.argv) > 3 else 20
if cname == "custom":     # QB_CFG=model,H,L,S,B,V
    mo, H_, L_, S_, B_, V_ = os.environ["QB_CFG"].split(",")
    BN.CONFIGS["custom"] = dict(model=mo, rnn_size=int(H_), num_layers=int(L_), seq_length=int(S_), batch_size=int(B_), vocab_size=int(V_), desc="custom")
c = BN.CONFIGS[cname]
margs = argparse.Namespace(grad_clip=5.0, output_keep_prob=1.0, input_keep_prob=1.0, **{k: v for k, v in c.items() if k != "desc"})
B, S = c["batch_size"], c["seq_length"]
X, Y = BN.synthetic_ids(c, K + 3, 1234)
Xd, Yd = torch.from_numpy(X).cuda(), torch.from_numpy(Y).cuda()
for prec in precs:
    m = Model(margs, device="cuda:0", seed=0, precision=prec); m.assign_lr(0.002)
    losses = []
    for k in range(3):
        m.step_async(Xd[:, k*S:(k+1)*S], Yd[:, k*S:(k+1)*S]); losses.append(float(m.grads[m.P+1]))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(K):
        m.step_async(Xd[:, (3+k)*S:(4+k)*S], Yd[:, (3+k)*S:(4+k)*S])
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / K
    import ctypes
    from distributed_char_rnn_b200 import _lib
    lib = _lib.load(); arr = (ctypes.c_double * _lib.NUM_PHASES)(); kms = (ctypes.c_double * 2)(); kln = (ctypes.c_int64 * 2)()
    lib.crnn_profile_enable(m._h, 1)
    for k in range(5):
        m.step_async(Xd[:, (3+k)*S:(4+k)*S], Yd[:, (3+k)*S:(4+k)*S]); lib.crnn_profile_read(m._h, arr, 0); lib.crnn_profile_read_kernels(m._h, kms, kln, 0)
    torch.cuda.synchronize(); lib.crnn_profile_enable(m._h, 0)
    ph = {lib.crnn_profile_phase_name(i).decode(): round(arr[i] / 5, 3) for i in range(_lib.NUM_PHASES) if arr[i] > 0}
    print("   phases", ph, "kernel fwd/bwd ms per launch %.4f %.4f (%d, %d launches/step)" % (kms[0] / max(1, kln[0]), kms[1] / max(1, kln[1]), kln[0] // 5, kln[1] // 5))
    print("QB %s %s env=%s ms/step %.4f Mchars/s %.3f losses %s final %.5f" % (cname, prec, {k: v for k, v in os.environ.items() if k.startswith("CRNN_")}, ms, B*S/ms/1e3, ["%.5f" % l for l in losses], float(m.grads[m.P+1])), flush=True)
    del m; torch.cuda.empty_cache()
