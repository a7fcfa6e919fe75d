"""Generates tests/golden/hash_vectors.json.

Sources of truth (none of them is this repository's code):
  * the three known-answer values asserted by the reference's own tests
    (/root/reference/lib/tokens/src/lib.rs:539-545, :604, :621; lib/llm/src/tokens.rs:375-376), copied as literals;
  * python `xxhash` (XXH 0.8.2), an independent implementation of the frozen XXH3-64 algorithm, for every other
    length class (the reference's tests only pin the 9-16 byte class).
Run: python tests/golden/make_hash_vectors.py   (needs only numpy + xxhash; no GPU, no reference tree)
"""
import json
import os
import struct

import numpy as np
import xxhash

HERE = os.path.dirname(os.path.abspath(__file__))


def h(data: bytes, seed: int) -> int:
    return xxhash.xxh3_64_intdigest(data, seed=seed)


def tok_bytes(tokens) -> bytes:
    return b"".join(struct.pack("<I", int(t) & 0xFFFFFFFF) for t in tokens)


def split(tokens, bs, salt):
    out = []
    parent = None
    for b in range(len(tokens) // bs):
        bh = h(tok_bytes(tokens[b * bs:(b + 1) * bs]), salt)
        sh = bh if parent is None else h(struct.pack("<QQ", parent, bh), salt)
        out.append([str(bh), str(sh)])
        parent = sh
    return out


def main():
    rng = np.random.Generator(np.random.PCG64(20250101))
    doc = {
        "reference_known_answers": {
            "comment": "lib/tokens/src/lib.rs:539-545: tokens 1..=10, block_size 4, salt 1337",
            "tokens": list(range(1, 11)), "block_size": 4, "salt": 1337,
            "block0_hash_and_seq": "14643705804678351452",
            "block1_hash": "16777012769546811212",
            "block1_seq": "4945711292740353085",
        },
        "raw": [],      # (hex bytes, seed) -> digest, one per length class edge
        "sequences": [],
    }
    lens = [0, 1, 2, 3, 4, 5, 8, 9, 15, 16, 17, 31, 32, 33, 44, 64, 96, 127, 128, 129, 160, 192, 239, 240, 241, 255,
            256, 257, 320, 448, 512, 1023, 1024, 1025, 2048, 4096]
    for L in lens:
        data = bytes(rng.integers(0, 256, size=L, dtype=np.uint8))
        for seed in (0, 1337, 0xDEADBEEFCAFEF00D, (1 << 64) - 1):
            doc["raw"].append({"hex": data.hex(), "seed": str(seed), "digest": str(h(data, seed))})
    cases = [(list(range(192)), 64, 1337), (list(range(192)), 64, 0), (list(range(1, 11)), 4, 1337),
             ([-1, -2, 5, 2 ** 31 - 1, 7, 8, 9, 10, 11], 4, 42)]
    for bs in (4, 11, 16, 32, 64, 128):
        n = 3 * bs + bs // 2
        toks = [int(x) for x in rng.integers(0, 128256, size=n)]
        cases.append((toks, bs, int(rng.integers(0, 2 ** 63))))
    for toks, bs, salt in cases:
        doc["sequences"].append({"tokens": toks, "block_size": bs, "salt": str(salt), "blocks": split(toks, bs, salt)})
    # the in-crate twin's first-block push_token quirk (lib/llm/src/tokens.rs:131-134): hash([0, block_hash])
    bh = h(tok_bytes([1, 2, 3, 4]), 1337)
    doc["twin_first_block_push"] = {"tokens": [1, 2, 3, 4], "seq": str(h(struct.pack("<QQ", 0, bh), 1337))}
    with open(os.path.join(HERE, "hash_vectors.json"), "w") as f:
        json.dump(doc, f, indent=1)
    print("wrote", len(doc["raw"]), "raw and", len(doc["sequences"]), "sequence vectors")


if __name__ == "__main__":
    main()
