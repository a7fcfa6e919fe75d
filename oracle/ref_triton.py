"""oracle/ref_triton.py — TEST INFRASTRUCTURE, never imported by the product (dynamo_exp_b200/).

Runs the REFERENCE'S OWN Triton rearrange kernels (`rearrange_tensors`, the new file
`vllm/distributed/device_communicators/kv_rearrange.py` inside
/root/reference/container/deps/vllm/vllm_v0.8.4-dynamo-kv-disagg-patch.patch:908-1036) so that the TP head-reshard
rows are pinned to reference code and not to a restatement written by the same hand as the
kernel. Nothing of the reference is copied into the repository:

  * `extract()` cuts the added file out of the patch where it lies under /root/reference and writes it to
    oracle/_ref/kv_rearrange.py. oracle/_ref/ is git-ignored build output (like oracle/_ref/libblock_copy_ref.so); it
    travels to the GPU box with the snapshot because /root/reference does not exist there.
  * `load()` imports that file. With a CUDA device the kernels are compiled by Triton and run on the GPU; without one
    (this container) they run under Triton's interpreter (TRITON_INTERPRET=1), which executes the same Python source
    with numpy — that is how tests/golden/rearrange_vectors.npz was produced (tests/golden/make_rearrange_vectors.py).
  * `python -m oracle.ref_triton --interpret in.npy out.npy d direction` is the subprocess form, because the interpreter
    switch must be set before `triton` is first imported.

On top of the kernel, `reference_write_blocks` / `reference_read_blocks` replay what the reference connector does with it
(patch:1296-1421): rearrange every (local range -> staging range) in place of the cache, then move the sub-block
descriptors `start*d + i*R + k` (patch:1215-1245) to / from the remote ranks' blocks; NIXL's transfer of a descriptor is
a byte copy, done here with tensor slicing.
"""
from __future__ import annotations

import importlib.util
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("KVDP_REFERENCE_ROOT", "/root/reference")
PATCH = os.path.join(REF_ROOT, "container", "deps", "vllm", "vllm_v0.8.4-dynamo-kv-disagg-patch.patch")
OUT = os.path.join(HERE, "_ref", "kv_rearrange.py")
_TARGET = "vllm/distributed/device_communicators/kv_rearrange.py"


def extract(force: bool = False):
    """Write oracle/_ref/kv_rearrange.py from the patch. Returns its path, or None when the reference tree is absent
    (then a previously extracted file, if any, is kept)."""
    if not os.path.exists(PATCH):
        return OUT if os.path.exists(OUT) else None
    if os.path.exists(OUT) and not force and os.path.getmtime(OUT) >= os.path.getmtime(PATCH):
        return OUT
    body, on = [], False
    with open(PATCH, "r", encoding="utf-8", errors="replace") as f:
        for line in f:
            line = line.rstrip("\n")
            if line.startswith("diff --git "):
                on = line.endswith(_TARGET)
                continue
            if on and line.startswith("+") and not line.startswith("+++"):
                body.append(line[1:])
    if not body:
        raise RuntimeError("the patch no longer adds " + _TARGET)
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    with open(OUT, "w") as f:
        f.write("\n".join(body) + "\n")
    return OUT


def available() -> bool:
    return os.path.exists(OUT)


_mod = None


def load():
    """Import oracle/_ref/kv_rearrange.py. Call only in a process where TRITON_INTERPRET is already what you want."""
    global _mod
    if _mod is None:
        if not available():
            raise RuntimeError("oracle/_ref/kv_rearrange.py is missing: run oracle.ref_triton.extract() where "
                               "/root/reference exists (build() does)")
        spec = importlib.util.spec_from_file_location("kvdp_ref_kv_rearrange", OUT)
        _mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_mod)
    return _mod


def rearrange_interpreted(t1: np.ndarray, t2: np.ndarray, d: int, direction: str):
    """The reference kernel on the CPU through Triton's interpreter, in a subprocess. Returns (t1, t2) after the call
    (write fills t2 from t1, read fills t1 from t2)."""
    import tempfile
    with tempfile.TemporaryDirectory() as tmp:
        a, b = os.path.join(tmp, "t1.npy"), os.path.join(tmp, "t2.npy")
        np.save(a, t1)
        np.save(b, t2)
        env = dict(os.environ, TRITON_INTERPRET="1", CUDA_VISIBLE_DEVICES="")
        r = subprocess.run([sys.executable, "-m", "oracle.ref_triton", "--interpret", a, b, str(d), direction],
                           cwd=os.path.dirname(HERE), env=env, capture_output=True, text=True, timeout=1200)
        if r.returncode != 0:
            raise RuntimeError("reference Triton kernel (interpreter) failed:\n" + r.stdout + r.stderr)
        return np.load(a), np.load(b)


# ---------------------------------------------------------------------------------- the connector's use of the kernel
def _ranges(ids):
    out = []
    for i, b in enumerate(ids):
        if i == 0 or b != ids[i - 1] + 1:
            out.append([b, b])
        else:
            out[-1][1] = b
    return out


def _same_length(src, dst):
    src, dst = [list(r) for r in src], [list(r) for r in dst]
    o_s, o_d, i, j = [], [], 0, 0
    while i < len(src) and j < len(dst):
        ls, ld = src[i][1] - src[i][0] + 1, dst[j][1] - dst[j][0] + 1
        n = min(ls, ld)
        o_s.append([src[i][0], src[i][0] + n - 1])
        o_d.append([dst[j][0], dst[j][0] + n - 1])
        if ls == n:
            i += 1
        else:
            src[i][0] += n
        if ld == n:
            j += 1
        else:
            dst[j][0] += n
    return o_s, o_d


def _sub_block(cache_kv, staging_ranges, block_id, i, d):
    """View of sub-block descriptor (block_id, i) of a [N, B, H, C] cache half: descriptor index start*d + i*R +
    (block_id - start) in units of block_len/d bytes (patch:1215-1245)."""
    N = cache_kv.shape[0]
    sub = cache_kv.reshape(N * d, -1)
    for lo, hi in staging_ranges:
        if lo <= block_id <= hi:
            return sub[lo * d + i * (hi - lo + 1) + (block_id - lo)]
    raise KeyError(block_id)


def reference_write_blocks(rearrange, local_caches, remote_caches_by_group, local_ids, staging_ids, remote_ids, d):
    """The data movement of the reference's write_blocks (patch:1361-1421) for one prefill rank whose decode engine has d
    times its TP degree. local_caches: per layer a torch tensor [2, N, B, H, C] (modified: staging blocks are overwritten,
    as in the reference); remote_caches_by_group[i]: per layer [2, Nr, B, H/d, C] of decode rank rank*d+i."""
    remote_ids = list(remote_ids)[:len(local_ids)]
    l_r, s_r = _same_length(_ranges(list(local_ids)), _ranges(list(staging_ids)))
    for (l0, l1), (s0, s1) in zip(l_r, s_r):
        for kv_cache in local_caches:
            for cache in kv_cache:
                rearrange(cache[l0:l1 + 1], cache[s0:s1 + 1], d, "write")
    for layer, kv_cache in enumerate(local_caches):
        for e in range(2):
            for i in range(d):
                dst = remote_caches_by_group[i][layer][e]
                dst2 = dst.reshape(dst.shape[0], -1)
                for sb, rb in zip(staging_ids, remote_ids):
                    dst2[rb] = _sub_block(kv_cache[e], s_r, sb, i, d).reshape(dst2[rb].shape)


def reference_read_blocks(rearrange, local_caches, remote_caches_by_group, local_ids, staging_ids, remote_ids, d):
    """The data movement of the reference's read_blocks (patch:1296-1359): fetch the sub-blocks of the d remote ranks
    into the staging range, then rearrange "read" staging -> local."""
    l_r, s_r = _same_length(_ranges(list(local_ids)), _ranges(list(staging_ids)))
    for layer, kv_cache in enumerate(local_caches):
        for e in range(2):
            for i in range(d):
                src = remote_caches_by_group[i][layer][e]
                src2 = src.reshape(src.shape[0], -1)
                for sb, rb in zip(staging_ids, remote_ids):
                    view = _sub_block(kv_cache[e], s_r, sb, i, d)
                    view.copy_(src2[rb].reshape(view.shape))
    for (l0, l1), (s0, s1) in zip(l_r, s_r):
        for kv_cache in local_caches:
            for cache in kv_cache:
                rearrange(cache[l0:l1 + 1], cache[s0:s1 + 1], d, "read")


def _main(argv):
    if len(argv) != 6 or argv[1] != "--interpret":
        print("usage: python -m oracle.ref_triton --interpret t1.npy t2.npy d write|read", file=sys.stderr)
        return 2
    os.environ["TRITON_INTERPRET"] = "1"
    import torch
    mod = load()
    t1, t2 = torch.from_numpy(np.load(argv[2])), torch.from_numpy(np.load(argv[3]))
    mod.rearrange_tensors(t1, t2, int(argv[4]), argv[5])
    np.save(argv[2], t1.numpy())
    np.save(argv[3], t2.numpy())
    return 0


if __name__ == "__main__":
    sys.exit(_main(sys.argv))
