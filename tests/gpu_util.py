"""helpers of the -m gpu parity tests: move seeded inputs to the device and bring results back in the
oracle's record layouts"""
import numpy as np
import torch

import nvbio_gpl_b200 as nvb
import pyoracle as po


def dev():
    return torch.device("cuda", 0)


def sset(sym, off):
    return nvb.StringSet.from_numpy(sym, off, dev())


def aligner(kind, ty, p):
    return nvb.Aligner(kind, ty, *p)


def sinks_np(t):
    a = t.cpu().numpy()
    out = np.zeros(a.shape[0], dtype=po.SINK_DT)
    out["score"] = a[:, 0]
    out["x"] = a[:, 1].view(np.uint32)
    out["y"] = a[:, 2].view(np.uint32)
    return out


def aln_np(t):
    a = t.cpu().numpy()
    out = np.zeros(a.shape[0], dtype=po.ALN_DT)
    out["score"] = a[:, 0]
    for k, name in enumerate(("src_x", "src_y", "sink_x", "sink_y")):
        out[name] = a[:, k + 1].view(np.uint32)
    return out


def tb_np(res):
    out, ops, n_ops = res
    return aln_np(out), ops.cpu().numpy(), n_ops.cpu().numpy().view(np.uint32)


def tb_equal(got, want):
    """(alignments, ops, n_ops) equality; ops compared on the valid prefix (strides may differ)"""
    (o1, ops1, n1), (o2, ops2, n2) = got, want
    if not ((o1 == o2).all() and (n1 == n2).all()):
        bad = np.nonzero(~((o1 == o2) & (n1 == n2)))[0][:5]
        print("mismatch at", bad, o1[bad], o2[bad], n1[bad], n2[bad])
        return False
    for i in range(len(n1)):
        k = min(int(n1[i]), ops1.shape[1], ops2.shape[1])
        if not (ops1[i, :k] == ops2[i, :k]).all():
            print("ops mismatch at", i)
            return False
    return True


def fm_index_pair(oracle, text):
    """(device FMIndex built by the product, oracle Index over the same host image)"""
    fmi = nvb.build_index(torch.from_numpy(text).to(dev()))
    bwt_occ, ssa = fmi.host_arrays()
    oidx = po.Index(fmi.length, fmi.primary, np.array(fmi.L2, dtype=np.uint32), bwt_occ, ssa)
    return fmi, oidx
