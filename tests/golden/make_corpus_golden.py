"""Golden INPUT of configs 1 and 2: the reference's own corpus, packed so it can travel to the GPU box.

``/root/reference`` does not exist there, and SURVEY.md §8(d) states config 1 / config 2 parity on
``data/tinyshakespeare/input.txt`` (1,115,394 bytes of ASCII, sha256 86c4e6aa9db7c042...).  This script
copies that data file byte for byte into ``tests/golden/tinyshakespeare.txt.xz`` (lzma, ~340 KB) and
records its digest; ``tests/test_gpu_configs.py`` unpacks it and checks the digest before use.

Run here (the reference tree is read-only and only present in the build container):
    python tests/golden/make_corpus_golden.py
"""
import hashlib
import json
import lzma
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/data/tinyshakespeare/input.txt"


def main():
    raw = open(SRC, "rb").read()
    with open(os.path.join(HERE, "tinyshakespeare.txt.xz"), "wb") as f:
        f.write(lzma.compress(raw, preset=9 | lzma.PRESET_EXTREME))
    meta = {"source": "data/tinyshakespeare/input.txt", "bytes": len(raw), "sha256": hashlib.sha256(raw).hexdigest()}
    json.dump(meta, open(os.path.join(HERE, "tinyshakespeare.json"), "w"), indent=1)
    print(meta)


if __name__ == "__main__":
    main()
