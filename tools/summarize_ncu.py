"""Condense ncu outputs (launch list csv, --page raw csv) into the markdown summaries kept under profiles/."""
import collections, csv, sys


def launches(path, out, title):
    rows = list(csv.reader(l for l in open(path) if not l.startswith("==")))
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        if len(r) <= vi:
            continue
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        a = agg.setdefault(r[ki].split("(")[0][:80], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    with open(out, "w") as f:
        f.write("# %s\n\n`ncu --metrics gpu__time_duration.sum --clock-control none` launch list, summed per kernel "
                "(cold-cache, serialised: compare shares, not absolutes).\n\n" % title)
        f.write("total %d launches, %.3f ms\n\n| kernel | launches | total us | share |\n|---|---|---|---|\n" % (len(rows) - 1, tot / 1e6))
        for n, a in sorted(agg.items(), key=lambda x: -x[1][1]):
            if a[1] / tot < 0.0005:
                continue
            f.write("| `%s` | %d | %.1f | %.1f%% |\n" % (n, a[0], a[1] / 1e3, 100 * a[1] / tot))


WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.sum",
        "l1tex__t_set_accesses_pipe_lsu_mem_global_op_red.sum", "lts__t_sectors_op_red.sum"]


def raw(path, out, title):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    with open(out, "w") as f:
        f.write("# %s\n\n`ncu --set full --clock-control none --import-source on`, one launch per kernel, selected raw metrics.\n" % title)
        for r in rows[2:]:
            f.write("\n## %s\n\n| metric | value | unit |\n|---|---|---|\n" % r[hdr.index("Kernel Name")].split("(")[0])
            for w in WANT:
                if w in hdr:
                    i = hdr.index(w)
                    f.write("| %s | %s | %s |\n" % (w, r[i], units[i]))


def main():
    kind, path, out, title = sys.argv[1:5]
    {"launches": launches, "raw": raw, "traffic": traffic}[kind](path, out, title)


def traffic(path, out, title):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of each of our kernels -> JSON read by bench.py
    (the `traffic` key of the roofline object)."""
    import json, re
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    ir, iw, ik = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name")
    mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    names = {"render_fwd_kernel": "render_fwd", "render_bwd_kernel": "render_bwd", "mlp_fwd_kernel": "mlp_fwd",
             "mlp_dgrad_kernel": "mlp_dgrad", "mlp_wgrad_kernel": "mlp_wgrad", "preprocess_fwd_kernel": "preprocess_fwd",
             "preprocess_bwd_kernel": "preprocess_bwd", "duplicate_with_keys_kernel": "duplicate_with_keys",
             "DeviceRadixSortOnesweepKernel": "radix_sort_pass", "tile_sort_kernel": "tile_sort",
             "tile_count_kernel": "tile_count", "tile_scatter_kernel": "tile_scatter"}
    acc = {}
    for r in rows[2:]:
        for pat, short in names.items():
            if re.search(pat, r[ik]):
                b = float(r[ir].replace(",", "")) * mult[units[ir]] + float(r[iw].replace(",", "")) * mult[units[iw]]
                acc.setdefault(short, []).append(b)
    res = {"source": title, "unit": "bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum)",
           "kernels": {k: sum(v) / len(v) for k, v in acc.items()}}
    if "radix_sort_pass" in res["kernels"]:
        res["kernels"]["radix_sort"] = res["kernels"]["radix_sort_pass"] * 6  # 6 onesweep passes per sort at C2 (44 key bits)
    json.dump(res, open(out, "w"), indent=1)


if __name__ == "__main__":
    main()
