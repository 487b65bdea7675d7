"""Generate the integer golden vectors of the data contract from the REFERENCE's own code.

Run in the build container (needs /root/reference):  python tests/golden/make_data_golden.py
Imports the reference's ``utils.py`` and runs its ``data_splitter.py`` as a subprocess on
small synthetic corpora written here (ASCII with count ties, multi-byte UTF-8, CR/LF), then
stores what they produced in ``data_golden.json``.  The tests replay the same corpora through
``distributed_char_rnn_b200.utils`` on any box (the reference is not present on the GPU box).
"""
import hashlib
import importlib.util
import json
import os
import pickle
import shutil
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def corpora():
    rng = np.random.Generator(np.random.PCG64(7))
    ascii_alpha = "etaoin shrdlu\n"
    a = "".join(ascii_alpha[i] for i in rng.integers(0, len(ascii_alpha), 4000))
    ties = "abcabcabc zzz yyy xxx\r\nq"                       # many equal counts, CR survives
    uni = "".join(rng.choice(list(u"añ€😀ßz \n"), 3000))       # 1-4 byte UTF-8
    return {"ascii": a * 3, "ties": ties * 40, "unicode": uni}


def main():
    spec = importlib.util.spec_from_file_location("ref_utils", os.path.join(REF, "utils.py"))
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    out = {"corpora": {}, "tinyshakespeare": {}}
    for name, text in corpora().items():
        d = tempfile.mkdtemp()
        with open(os.path.join(d, "input.txt"), "wb") as f:
            f.write(text.encode("utf-8"))
        B, S = 4, 7
        ld = ref.TextLoader(d, B, S)
        rec = {"text": text, "chars": list(ld.chars), "vocab_size": ld.vocab_size, "batch_size": B, "seq_length": S,
               "num_batches": ld.num_batches, "tensor_dtype": str(np.load(os.path.join(d, "data.npy")).dtype),
               "tensor": np.load(os.path.join(d, "data.npy")).tolist(),
               "x0": ld.x_batches[0].tolist(), "y0": ld.y_batches[0].tolist(),
               "x_last": ld.x_batches[-1].tolist(), "y_last": ld.y_batches[-1].tolist()}
        # reload path (utils.py:53-60)
        ld2 = ref.TextLoader(d, B, S)
        assert ld2.chars == ld.chars and (ld2.tensor == ld.tensor).all()
        # data_splitter.py as a script, on a writable copy
        n = len(np.load(os.path.join(d, "data.npy")))
        parts = [p for p in (2, 3, 4) if n % p == 0]
        rec["split_ok"] = parts
        rec["split_bad"] = [p for p in (2, 3, 4, 7) if n % p != 0][:1]
        rec["shards"] = {}
        for p in parts:
            od = os.path.join(d, "sh%d" % p)
            subprocess.check_call([sys.executable, os.path.join(REF, "data_splitter.py"), "--data_dir", d,
                                   "--num_parts", str(p), "--out_dir", od], stdout=subprocess.DEVNULL)
            rec["shards"][str(p)] = [hashlib.sha256(np.load(os.path.join(od, "data-%d.npy" % i)).astype("<i8").tobytes()).hexdigest()
                                     for i in range(p)]
        for p in rec["split_bad"]:
            r = subprocess.run([sys.executable, os.path.join(REF, "data_splitter.py"), "--data_dir", d, "--num_parts", str(p),
                                "--out_dir", os.path.join(d, "bad")], capture_output=True, text=True)
            assert r.returncode != 0 and "ValueError" in r.stderr
        out["corpora"][name] = rec
        shutil.rmtree(d)
    # tinyshakespeare known answers: hashes only, the corpus is not copied
    d = tempfile.mkdtemp()
    shutil.copy(os.path.join(REF, "data/tinyshakespeare/input.txt"), d)
    ld = ref.TextLoader(d, 50, 50)
    full = np.load(os.path.join(d, "data.npy"))
    ts = out["tinyshakespeare"]
    ts["input_sha256"] = hashlib.sha256(open(os.path.join(d, "input.txt"), "rb").read()).hexdigest()
    ts["chars"] = "".join(ld.chars)
    ts["tensor_i4_sha256"] = hashlib.sha256(full.astype("<i4").tobytes()).hexdigest()
    ts["num_batches"] = ld.num_batches
    ts["x0_row0"] = ld.x_batches[0][0, :5].tolist()
    ts["y0_row0"] = ld.y_batches[0][0, :5].tolist()
    ts["x1_row0"] = ld.x_batches[1][0, :5].tolist()
    ts["x0_row1"] = ld.x_batches[0][1, :5].tolist()
    ts["y_last_wrap"] = int(ld.y_batches[445][49, -1])
    sh = [np.load(os.path.join(REF, "sharded_data/data-%d.npy" % i)) for i in range(2)]
    ts["shipped_shard_dtype"] = str(sh[0].dtype)
    ts["shipped_shard_sha256"] = [hashlib.sha256(s.astype("<i4").tobytes()).hexdigest() for s in sh]
    ts["shards_concat_equal_full"] = bool((np.concatenate(sh) == full).all())
    ts["vocab_pkl_chars"] = "".join(pickle.load(open(os.path.join(d, "vocab.pkl"), "rb")))
    shutil.rmtree(d)
    with open(os.path.join(HERE, "data_golden.json"), "w") as f:
        json.dump(out, f, ensure_ascii=True, indent=0)
    print("wrote data_golden.json", os.path.getsize(os.path.join(HERE, "data_golden.json")))


if __name__ == "__main__":
    main()
