#!/usr/bin/env python
"""Regenerates tests/golden/rearrange_vectors.npz by RUNNING THE REFERENCE: the Triton kernels of
vllm_v0.8.4-dynamo-kv-disagg-patch.patch:908-1036 (extracted to oracle/_ref/kv_rearrange.py by oracle/ref_triton.py),
executed by Triton's interpreter on the CPU, plus the reference connector's descriptor arithmetic around them
(oracle.ref_triton.reference_write_blocks / reference_read_blocks, patch:1215-1245, 1296-1421).

    python tests/golden/make_rearrange_vectors.py [out.npz]        (needs /root/reference or a prior extract)

Inputs are not stored: they are `case_input(seed, shape)` of tests/golden/rearrange_cases.py, which the tests call too. Every shape keeps
N*B*H*C a multiple of 1024 because the reference kernels have no bounds mask.
"""
import os
import sys

os.environ["TRITON_INTERPRET"] = "1"          # before triton is imported anywhere
os.environ.setdefault("CUDA_VISIBLE_DEVICES", "")

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from tests.golden.rearrange_cases import (CONNECTOR_CASES, KERNEL_CASES, case_input,  # noqa: E402
                                          connector_inputs)


def main(out_path: str) -> None:
    import torch

    from oracle import ref_triton
    if ref_triton.extract() is None:
        raise SystemExit("the reference tree is absent and nothing was extracted before")
    ref = ref_triton.load()
    out = {}
    for i, (name, shape, d) in enumerate(KERNEL_CASES):
        t1 = torch.from_numpy(case_input(10 + i, shape))
        t2 = torch.zeros_like(t1)
        ref.rearrange_tensors(t1, t2, d, "write")
        out[name + "_write"] = t2.numpy().copy()
        t2r = torch.from_numpy(case_input(50 + i, shape))
        t1r = torch.zeros_like(t2r)
        ref.rearrange_tensors(t1r, t2r, d, "read")
        out[name + "_read"] = t1r.numpy().copy()
    for i, (name, layers, nl, nr, B, H, C, d, lid, sid, rid) in enumerate(CONNECTOR_CASES):
        local, remote = connector_inputs(i, layers, nl, nr, B, H, C, d)
        lt = [torch.from_numpy(a.copy()) for a in local]
        rt = [[torch.from_numpy(a.copy()) for a in grp] for grp in remote]
        ref_triton.reference_write_blocks(ref.rearrange_tensors, lt, rt, lid, sid, rid, d)
        for g in range(d):
            for l in range(layers):
                out["%s_write_g%d_l%d" % (name, g, l)] = rt[g][l].numpy().copy()
        lt = [torch.from_numpy(a.copy()) for a in local]
        rt = [[torch.from_numpy(a.copy()) for a in grp] for grp in remote]
        ref_triton.reference_read_blocks(ref.rearrange_tensors, lt, rt, lid, sid, rid, d)
        for l in range(layers):
            # only the local (non-staging) blocks are the result; staging holds the reference's intermediate
            out["%s_read_l%d" % (name, l)] = lt[l].numpy()[:, lid].copy()
    np.savez_compressed(out_path, **out)
    print("wrote", out_path, "with", len(out), "arrays,", os.path.getsize(out_path), "bytes")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.abspath(__file__)),
                                                             "rearrange_vectors.npz"))
