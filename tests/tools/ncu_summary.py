"""Summarise an `ncu --set full --import-source on` capture (read here, no GPU needed) into a text block for profiles/.

    python tests/tools/ncu_summary.py gpurun_out/prof_decode.ncu-rep [--launch 0] > profiles/rNN_ncu_<kernel>.txt
Prints: duration, DRAM bytes (the `traffic` figure of bench.py's roofline object), DRAM %, pipe utilisation, registers,
the warp-stall histogram of the source page, the SASS opcode mix and the hottest instructions.
"""
import collections
import csv
import io
import re
import subprocess
import sys


def ncu_csv(rep, page):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    launch = int(sys.argv[sys.argv.index("--launch") + 1]) if "--launch" in sys.argv else 0
    raw = ncu_csv(rep, "raw")
    h, units, rows = raw[0], raw[1], raw[2:]
    col = {n: i for i, n in enumerate(h)}
    r = rows[launch]
    print(f"# {rep}: launch {launch} of {len(rows)}: {r[col['Kernel Name']][:100]}")
    print(f"grid {r[col['Grid Size']]} block {r[col['Block Size']]}")
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum.per_second",
            "lts__t_sector_hit_rate.pct", "launch__registers_per_thread", "launch__occupancy_limit_shared_mem",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
            "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
            "sm__cycles_elapsed.max", "smsp__inst_executed.sum",
            "lts__t_bytes.sum", "lts__t_sectors_srcunit_tex_op_read.sum", "l1tex__m_xbar2l1tex_read_bytes.sum",
            "l1tex__m_xbar2l1tex_read_bytes.sum.per_second", "lts__t_bytes.sum.per_second",
            "smsp__inst_executed_op_tma.sum", "sm__inst_executed_pipe_tma.sum", "l1tex__m_l1tex2xbar_write_bytes.sum",
            "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
            "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"]
    for w in want:
        if w in col:
            print(f"{w:75s} {r[col[w]]:>16s} {units[col[w]]}")
    try:
        rd = float(r[col["dram__bytes_read.sum"]].replace(",", "")); wr = float(r[col["dram__bytes_write.sum"]].replace(",", ""))
        mul = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}
        t = rd * mul[units[col["dram__bytes_read.sum"]]] + wr * mul[units[col["dram__bytes_write.sum"]]]
        print(f"traffic (dram read + write) per launch: {t:.0f} bytes")
    except Exception as e:  # noqa: BLE001
        print("traffic: n/a", e)

    src = ncu_csv(rep, "source")
    secs = [i for i, x in enumerate(src) if x and x[0] == "Kernel Name"]
    if not secs:
        return
    s = secs[min(launch, len(secs) - 1)]
    e = secs[secs.index(s) + 1] if secs.index(s) + 1 < len(secs) else len(src)
    hh, body = src[s + 1], [x for x in src[s + 2:e] if len(x) > 10]
    ci = {n: i for i, n in enumerate(hh)}
    tot = sum(int(x[ci["# Samples"]]) for x in body)
    stalls = [n for n in hh if n.startswith("stall_") and "Not Issued" not in n]
    print(f"\n## warp-stall samples (all): {tot} over {len(body)} SASS instructions")
    agg = {n: sum(int(x[ci[n]]) for x in body) for n in stalls}
    for n, v in sorted(agg.items(), key=lambda kv: -kv[1]):
        if v:
            print(f"  {n:26s}{v:8d} {100 * v / max(tot, 1):5.1f}%")
    cnt = collections.Counter()
    ninst = 0
    for x in body:
        n = int(x[ci["Instructions Executed"]]); ninst += n
        op = re.sub(r"^@!?U?P\d+\s+", "", x[1].strip()).split()[0].split(".")[0]
        cnt[op] += n
    print(f"\n## SASS opcode mix (warp instructions executed: {ninst})")
    for k, v in cnt.most_common(24):
        print(f"  {k:12s}{v:10d} {100 * v / max(ninst, 1):5.1f}%")
    key = ["UTCQMMA", "UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UBLKCP", "SYNCS", "HMMA", "STAS", "MUFU", "F2FP"]
    print("\n## tensor-core / TMEM / TMA instructions executed (warp instructions): " + ", ".join(f"{k} {cnt[k]}" for k in key if cnt.get(k)))
    print("\n## hottest instructions (samples, SASS, top stall reasons)")
    for x in sorted(body, key=lambda x: -int(x[ci["# Samples"]]))[:24]:
        st = {n[6:]: int(x[ci[n]]) for n in stalls if int(x[ci[n]]) > 0}
        top = ", ".join(f"{k}={v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"  {x[ci['# Samples']]:>5s}  {x[1].strip()[:64]:64s} {top}")


if __name__ == "__main__":
    main()
