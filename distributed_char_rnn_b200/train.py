"""``train.py`` of the reference, on the B200 engine: same flags, files and loop semantics.

Reference behaviour kept (``train.py:73-218``): loader -> ``--init_from`` compatibility asserts ->
``config.pkl`` / ``chars_vocab.pkl`` -> model -> per epoch: ``lr = learning_rate * decay_rate**e``,
reset the batch pointer, zero the RNN state; per batch: feed ``x, y`` and the previous final
state, run one step, print ``"{step}/{total} (epoch {e}), train_loss = ..., time/batch = ..."``;
save at ``(e*num_batches+b) % save_every == 0`` and at the very last step.

What changes (SURVEY §8 e): ``--distributed`` no longer builds a parameter server.  Every
``--job_name worker`` process is one synchronous data-parallel replica on its own GPU with its own
``data_splitter`` shard (``--tensor_file``); ranks exchange ONE all-reduce of the flat gradient
buffer per step over NCCL.  ``--job_name ps`` has nothing to do and exits 0.  World size and
rendezvous come from ``--worker_hosts`` (first entry = master address:port) unless the launcher
already set ``WORLD_SIZE`` / ``MASTER_ADDR`` (torchrun).  Reference defects D1-D5 are
not reproduced: single mode runs, checkpoints are written, ``--model gru|rnn`` carry their state.
"""
from __future__ import print_function

import argparse
import os
import pickle
import time

import numpy as np

from .utils import TextLoader
from . import checkpoint as ckpt_io


def build_parser():
    p = argparse.ArgumentParser(formatter_class=argparse.ArgumentDefaultsHelpFormatter)
    # data and checkpoint directories (train.py:17-35)
    p.add_argument('--data_dir', type=str, default='data/tinyshakespeare',
                   help='data directory containing input.txt with training examples')
    p.add_argument('--save_dir', type=str, default='save', help='directory to store checkpointed models')
    p.add_argument('--log_dir', type=str, default='logs', help='directory to store tensorboard logs')
    p.add_argument('--save_every', type=int, default=1000,
                   help='Save frequency. Number of passes between checkpoints of the model.')
    p.add_argument('--init_from', type=str, default=None,
                   help="continue training from saved model at this path (must contain config.pkl, "
                        "chars_vocab.pkl, checkpoint and model.ckpt-*; model, rnn_size, num_layers and "
                        "seq_length must match)")
    # model (train.py:37-42)
    p.add_argument('--model', type=str, default='lstm', help='lstm, rnn, gru, or nas')
    p.add_argument('--rnn_size', type=int, default=128, help='size of RNN hidden state')
    p.add_argument('--num_layers', type=int, default=2, help='number of layers in the RNN')
    # optimisation (train.py:44-61)
    p.add_argument('--seq_length', type=int, default=50, help='RNN sequence length. Number of timesteps to unroll for.')
    p.add_argument('--batch_size', type=int, default=50, help='minibatch size (per replica)')
    p.add_argument('--num_epochs', type=int, default=50, help='number of epochs')
    p.add_argument('--grad_clip', type=float, default=5., help='clip gradients at this value')
    p.add_argument('--learning_rate', type=float, default=0.002, help='learning rate')
    p.add_argument('--decay_rate', type=float, default=0.97, help='per-epoch learning-rate decay')
    p.add_argument('--output_keep_prob', type=float, default=1.0, help='probability of keeping weights in the hidden layer')
    p.add_argument('--input_keep_prob', type=float, default=1.0, help='probability of keeping weights in the input layer')
    # distributed (train.py:63-69)
    p.add_argument('--distributed', action='store_true', help="Indicates running in distributed mode")
    p.add_argument('--ps_hosts', default=None, help="PS HOSTS (accepted, unused: there is no parameter server)")
    p.add_argument('--worker_hosts', default=None, help="WORKER HOSTS: one data-parallel replica each")
    p.add_argument('--job_name', choices=['ps', 'worker'], default=None, help="Job name. Must be ps/worker")
    p.add_argument('--task_index', type=int, default=None, help="Index of task for given job (= replica rank)")
    p.add_argument('--tensor_file', default=None, help="Tensor file of the specific training data for given node.")
    # engine options (not in the reference)
    p.add_argument('--precision', choices=['auto', 'fp32', 'bf16x3', 'bf16'], default='auto',
                   help='bf16x3: tcgen05 contractions on split-bf16 operands, fp32-grade (1e-4 parity); fp32: FFMA contractions '
                        '(1e-4 parity); bf16: plain bf16 tcgen05 contractions (1e-2); auto: bf16x3 where the shape allows it, else fp32')
    p.add_argument('--seed', type=int, default=None, help='seed of the variable initialiser (reference: unseeded)')
    p.add_argument('--max_steps', type=int, default=None, help='stop after this many steps (smoke tests)')
    return p


def _init_distributed(args):
    """Map the reference's cluster flags onto one NCCL process group (SURVEY §8 b)."""
    import torch
    import torch.distributed as dist
    from .parallel import cluster_from_flags
    rank, world, addr, port = cluster_from_flags(args)
    os.environ.setdefault("MASTER_ADDR", addr)
    os.environ.setdefault("MASTER_PORT", str(port))
    local = int(os.environ.get("LOCAL_RANK", rank % max(1, torch.cuda.device_count())))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    return dist.group.WORLD, rank, world, local


def check_init_from(args, data_loader):
    """The reference's compatibility asserts, same messages (``train.py:88-107``)."""
    d = args.init_from
    assert os.path.isdir(d), " %s must be a a path" % d
    assert os.path.isfile(os.path.join(d, "config.pkl")), "config.pkl file does not exist in path %s" % d
    assert os.path.isfile(os.path.join(d, "chars_vocab.pkl")), "chars_vocab.pkl.pkl file does not exist in path %s" % d
    ckpt = ckpt_io.latest_checkpoint(d)
    assert ckpt, "No checkpoint found"
    with open(os.path.join(d, 'config.pkl'), 'rb') as f:
        saved_model_args = pickle.load(f)
    for checkme in ["model", "rnn_size", "num_layers", "seq_length"]:
        assert vars(saved_model_args)[checkme] == vars(args)[checkme], \
            "Command line argument and saved model disagree on '%s' " % checkme
    with open(os.path.join(d, 'chars_vocab.pkl'), 'rb') as f:
        saved_chars, saved_vocab = pickle.load(f)
    assert saved_chars == data_loader.chars, "Data and loaded model disagree on character set!"
    assert saved_vocab == data_loader.vocab, "Data and loaded model disagree on dictionary mappings!"
    return ckpt


def train(args):
    if args.distributed and args.job_name == "ps":
        print("job_name=ps: this engine has no parameter server (replicas all-reduce their gradients); nothing to do")
        return 0
    if args.distributed:
        if args.tensor_file is None:                      # train.py:79-80
            raise Exception("Tensor file must be provided.")
        data_loader = TextLoader(args.data_dir, args.batch_size, args.seq_length, tensor_file=args.tensor_file)
    else:
        data_loader = TextLoader(args.data_dir, args.batch_size, args.seq_length)
    args.vocab_size = data_loader.vocab_size

    ckpt = check_init_from(args, data_loader) if args.init_from is not None else None

    os.makedirs(args.save_dir, exist_ok=True)
    with open(os.path.join(args.save_dir, 'config.pkl'), 'wb') as f:
        pickle.dump(args, f)
    with open(os.path.join(args.save_dir, 'chars_vocab.pkl'), 'wb') as f:
        pickle.dump((data_loader.chars, data_loader.vocab), f)

    import torch
    from .model import Model
    pg, rank, world, local = (None, 0, 1, torch.cuda.current_device() if torch.cuda.is_available() else 0)
    if args.distributed:
        pg, rank, world, local = _init_distributed(args)
    model = Model(args, device="cuda:%d" % local, seed=args.seed, precision=args.precision, process_group=pg)
    if ckpt is not None:
        model.load_state_dict(ckpt_io.load(ckpt))
        model.sync_replicas()
    chief = rank == 0

    writer = None
    if chief:
        try:
            from torch.utils.tensorboard import SummaryWriter
            writer = SummaryWriter(os.path.join(args.log_dir, time.strftime("%Y-%m-%d-%H-%M-%S")))
        except Exception:
            writer = None

    # the whole shard lives on the device; a batch is a column block of it (utils.py:79-82)
    X, Y = data_loader.batch_matrices()
    Xd, Yd = torch.from_numpy(X).to(model.device), torch.from_numpy(Y).to(model.device)
    S, nb = args.seq_length, data_loader.num_batches
    if world > 1:                                         # replicas step in lock-step: same batch count
        t = torch.tensor([nb], device=model.device)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MIN, group=pg)
        nb = int(t.item())
    total = args.num_epochs * nb
    steps = 0
    for e in range(args.num_epochs):
        model.assign_lr(np.float32(args.learning_rate) * np.float32(args.decay_rate) ** np.float32(e))   # train.py:147-148,187
        data_loader.reset_batch_pointer()
        state = model.zero_state()                        # train.py:190
        for b in range(nb):
            start = time.time()
            train_loss, state = model.step(Xd[:, b * S:(b + 1) * S], Yd[:, b * S:(b + 1) * S], state)
            end = time.time()
            gs = e * nb + b
            if writer is not None and (world == 1 or gs % 100 == 0):          # train.py:162,202
                writer.add_scalar("train_loss", train_loss, gs)
            print("{}/{} (epoch {}), train_loss = {:.3f}, time/batch = {:.3f}"
                  .format(model.global_step, total, e, train_loss, end - start))
            last = (e == args.num_epochs - 1 and b == nb - 1)
            steps += 1
            stop = args.max_steps is not None and steps >= args.max_steps
            if chief and (gs % args.save_every == 0 or last or stop):        # train.py:210-217
                path = ckpt_io.save(args.save_dir, model.state_dict(), gs)
                print("model saved to {}".format(path))
            if stop:
                break
        if args.max_steps is not None and steps >= args.max_steps:
            break
    if writer is not None:
        writer.close()
    if world > 1:
        torch.distributed.destroy_process_group()
    return 0


def main(argv=None):
    return train(build_parser().parse_args(argv))


if __name__ == '__main__':
    main()
