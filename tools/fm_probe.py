"""tools/fm_probe.py -- tuning probe of nvb_fm_match: plain one-query-per-thread kernel vs the batch-sharing path
(fm_match_sorted.cu) over sort widths and run lengths, on the configs[1] index (100 Mbp, L2-resident) and the
configs[3] index (3.1 Gbp, DRAM-resident). Prints one JSON line per setting.  python tools/fm_probe.py [genome sizes...]"""
import ctypes
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (workload generators)
import nvbio_gpl_b200 as nvb  # noqa: E402
from nvbio_gpl_b200._lib import check, current_stream_ptr  # noqa: E402
from nvbio_gpl_b200.fmindex import _set_c  # noqa: E402


def main():
    sizes = [int(a) for a in sys.argv[1:]] or [100_000_000, 3_100_000_000]
    dev = torch.device("cuda", 0)
    ctx = nvb.Context.default(0)
    L = nvb.lib()
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    nq = int(os.environ.get("PROBE_QUERIES", 10_000_000))
    for G in sizes:
        t0 = time.time()
        genome = bench.gen_genome(G, 2, dev)
        fmi = nvb.build_index(genome)
        torch.cuda.synchronize()
        build_s = time.time() - t0
        qsym, qoff = bench.gen_queries(genome, nq, 100, dev)
        del genome
        torch.cuda.empty_cache()
        fc, qset = fmi.c(), _set_c(qsym, qoff)
        ranges = torch.empty((nq, 2), dtype=torch.int32, device=dev)

        def run():
            check(L.nvb_fm_match(ctx.h, ctypes.byref(fc), qset, ranges.data_ptr(), current_stream_ptr()), ctx.h)

        def timed(tag, **kw):
            ts = bench.time_steps(run, 5, 2, lambda: flush_buf.fill_(1))
            ms = sum(ts) / len(ts)
            print(json.dumps(dict(genome=G, queries=nq, tag=tag, ms=round(ms, 4), glookups_s=round(nq / ms / 1e6, 3), min_ms=round(min(ts), 4), build_s=round(build_s, 1), **kw)), flush=True)

        if os.environ.get("PROBE_ONLY") == "table":      # profiling aid: two calls of the default path, nothing else
            run(); run(); torch.cuda.synchronize()
            continue
        ctx.fm_batch_sharing(0xFFFFFFFF)
        timed("plain")
        want = ranges.clone()
        ctx.fm_batch_sharing(1 << 16)
        for k in os.environ.get("PROBE_K", "0").split(","):
            if k != "0":
                os.environ["NVB_FM_TABLE_K"] = k
            ranges.zero_()
            timed("batch_table", table_k=int(k))
            assert bool((ranges == want).all().item()), "batch table changed a range"
        os.environ.pop("NVB_FM_TABLE_K", None)
        if os.environ.get("PROBE_SORTED"):
            os.environ["NVB_FM_SHARE"] = "sort"
            for bits, r in ((24, 8), (24, 4), (32, 8)):
                os.environ["NVB_FM_SORT_BITS"], os.environ["NVB_FM_SORT_RUN"] = str(bits), str(r)
                ranges.zero_()
                timed("sorted", sort_bits=bits, run=r)
                assert bool((ranges == want).all().item()), "sorted path changed a range"
            for k in ("NVB_FM_SORT_BITS", "NVB_FM_SORT_RUN", "NVB_FM_SHARE"):
                os.environ.pop(k)
        del fmi, qsym, qoff, ranges, want
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
