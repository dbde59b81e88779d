"""Turn the outputs of tools/profile_pass.sh (gpurun_out/launches_*.csv, gpurun_out/prof_*_<kernel>.ncu-rep) into the
markdown summaries committed under profiles/.  Needs `ncu` on PATH (reads reports, no GPU).
usage: python tools/ncu_summary.py <tag> > profiles/<tag>_ncu.md          (tag e.g. r01b)"""
import collections
import csv
import io
import re
import subprocess
import sys

KERNELS = ["convnet_f16_kernel", "conv16_f16_kernel", "wgrad_f16_kernel", "wgrad_tc_kernel", "lincomb_kernel"]
METRICS = [
    ("launch__grid_size", "grid"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem / CTA"),
    ("gpu__time_duration.sum", "duration (cold cache, serialised)"),
    ("gpc__cycles_elapsed.max", "elapsed cycles"),
    ("sm__cycles_active.avg", "SM active cycles (avg over 148)"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active, % of active cycles"),
    ("sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor memory pipe active, % of active cycles"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput, % of peak (elapsed)"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy, %"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "shared-memory wavefronts (LSU)"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "shared-memory bank conflicts (LSU)"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM written"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput, % of peak"),
]


def raw(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return rows[0], rows[1], rows[2:]


def launches(path):
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg, tot = collections.OrderedDict(), 0.0
    for r in rows[1:]:
        name = re.sub(r"\(.*", "", r[ki])[:70]
        v = float(r[vi].replace(",", ""))
        v = v / 1000 if r[ui] == "ns" else v * 1000 if r[ui] == "ms" else v
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
        tot += v
    return agg, tot, len(rows) - 1


def main():
    global KERNELS
    tag = sys.argv[1]
    vid = len(sys.argv) > 2 and sys.argv[2] == "vidode"
    what = "one VidODE forward (configs[4]: 128x128x3, dopri5, B=32, 1 x B200)" if vid else \
        "one eager training step (configs[1]: rk4, B=128, 1 x B200)"
    script = "tools/vidode_one_forward.py" if vid else "tools/one_step.py"
    if vid:
        KERNELS = ["conv_tc_kernel128", "conv_tc_kernel64", "upsample2x_fwd_kernel", "conv_tc2_kernel"]
    print("# %s - ncu summary of %s\n" % (tag, what))
    print("Produced by `tools/profile_pass%s.sh` under gpurun (after the same script exited 0 without ncu), summarised by" % ("_vidode" if vid else ""))
    print("`tools/ncu_summary.py %s`.  Times under ncu are cold-cache and serialised: shares, not absolutes.\n" % " ".join(sys.argv[1:]))
    agg, tot, n = launches("gpurun_out/launches_%s.csv" % tag)
    print("## Launch list (`ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none python %s`)\n" % script)
    print("%d kernel launches, %.2f ms summed.\n" % (n, tot / 1000))
    print("| share | total us | launches | avg us | kernel |\n|---|---|---|---|---|")
    ours = 0.0
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        if "odex::" in k:
            ours += v[1]
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:24]:
        print("| %.1f %% | %.1f | %d | %.1f | `%s` |" % (100 * v[1] / tot, v[1], v[0], v[1] / v[0], k))
    print("\nlibodex kernels: %.1f %% of the summed device time; the rest is what is left on the torch side (%s).\n" % (
        100 * ours / tot, "BatchNorm / ReLU passes, the 3-channel stem, cat, layout copies, grid_sample" if vid else
        "first / last layers of ConvEncode / ConvDecode, activations, layout shuffles, the loss and Adam"))
    for k in KERNELS:
        try:
            hdr, units, rows = raw("gpurun_out/prof_%s_%s.ncu-rep" % (tag, k))
        except Exception as e:  # noqa: BLE001
            print("## %s: report missing (%s)\n" % (k, e))
            continue
        print("## `%s` (`ncu --set full --clock-control none --import-source on -k regex:%s`)\n" % (k, k))
        names = [r[hdr.index("Kernel Name")] for r in rows]
        print("| metric | " + " | ".join("`%s`" % re.sub(r"\(.*", "", nm).replace("void ", "")[:40] for nm in names) + " |")
        print("|---|" + "---|" * len(rows))
        for m, label in METRICS:
            if m not in hdr:
                continue
            i = hdr.index(m)
            print("| %s | " % label + " | ".join("%s %s" % (r[i], units[i]) for r in rows) + " |")
        if k == "conv_tc2_kernel" and not vid and any(", 1>" in nm for nm in names):
            print("\nThe instances captured here are the 1x1 `transform_z0` convolutions (`<N, 1>`: 6 MFLOP per sample, latency-bound "
                  "by construction); in this build the 3x3 64->64 shapes run inside `convnet_tc_kernel`, and the 3x3 `conv_tc2` "
                  "instances of the step (ConvEncode / ConvDecode helpers) are in the launch list above.")
        print()


if __name__ == "__main__":
    main()
