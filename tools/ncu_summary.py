"""Markdown summary of one `ncu --set full` capture (run where ncu is installed; no GPU needed):
python tools/ncu_summary.py <file.ncu-rep> "<title>" "<capture command>" > profiles/<name>.md
Prints the raw-page metrics the roofline discussion uses and, from the source page, how samples / executed
instructions / active lanes distribute over the code regions of the render kernel (found by landmarks in the SASS)."""
import csv, io, subprocess, sys

rep, title, cmd = sys.argv[1], sys.argv[2], sys.argv[3]
WANT = """dram__bytes_read.sum dram__bytes_write.sum gpu__time_duration.sum launch__block_size launch__grid_size
launch__registers_per_thread sm__cycles_active.avg sm__cycles_elapsed.avg sm__issue_active.avg.pct_of_peak_sustained_elapsed
smsp__inst_executed.sum smsp__thread_inst_executed_per_inst_executed.ratio smsp__warps_active.avg.per_cycle_active
smsp__warps_eligible.avg.per_cycle_active sm__warps_active.avg.pct_of_peak_sustained_active l1tex__t_sector_hit_rate.pct
lts__t_sector_hit_rate.pct l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed
sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active
smsp__sass_inst_executed_op_local_ld.sum smsp__sass_inst_executed_op_local_st.sum""".split()


def page(name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


raw = page("raw")
hdr, units, val = raw[0], raw[1], raw[2]
kname = val[hdr.index("Kernel Name")]
print(f"# {title}\n\nCapture: `{cmd}`\n\nKernel: `{kname}`\n\n| metric | value | unit |\n|---|---|---|")
for k, h in enumerate(hdr):
    if h in WANT or ("issue_stalled" in h and h.endswith("per_issue_active.ratio")):
        print(f"| {h} | {val[k]} | {units[k]} |")

src = page("source")
h2, data = src[1], src[2:]
iS, iN, iX, iT = h2.index("Source"), h2.index("# Samples"), h2.index("Instructions Executed"), h2.index("Thread Instructions Executed")
iNI = h2.index("stall_no_inst") if "stall_no_inst" in h2 else None
tot = sum(int(r[iN]) for r in data) or 1
totx = sum(int(r[iX]) for r in data) or 1
wide = [k for k, r in enumerate(data) if "LDG.E.ENL2.256" in r[iS]]
print(f"\n## Where the time goes (source page: {len(data)} SASS instructions, {tot} samples, {totx / 1e6:.0f} M warp instructions)\n")
# the first run of four consecutive 256-bit loads is the 4-wide node fetch; the loop runs from the vote before it
# to the next vote after the push/pop code
runs = [k for k in wide if all(k + j in wide for j in range(4))]
if runs:
    n0 = runs[0]
    head = max(k for k in range(n0) if "VOTE" in data[k][iS])
    tail = min(k for k in range(n0 + 4, len(data)) if "VOTE" in data[k][iS]) + 2
    regions = [("path logic, ray start, work fetch (before the loop)", 0, head - 6), ("wide traversal loop (node step + leaves)", head - 6, tail),
               ("d.x == 0 loop and epilogue", tail, len(data))]
    print("| region | SASS range | samples | warp instructions | lanes per instruction | no_instruction share of region samples |\n|---|---|---|---|---|---|")
    for name, a, b in regions:
        ch = data[a:b]
        sm = sum(int(r[iN]) for r in ch); ex = sum(int(r[iX]) for r in ch); th = sum(int(r[iT]) for r in ch)
        ni = sum(int(r[iNI]) for r in ch) if iNI is not None else 0
        print(f"| {name} | {a}-{b} | {100 * sm / tot:.1f} % | {100 * ex / totx:.1f} % ({ex / 1e6:.0f} M) | {th / max(ex, 1):.1f} | {100 * ni / max(sm, 1):.0f} % |")
    loop = data[head - 6:tail]
    lo = [r for r in loop if int(r[iX]) and int(r[iT]) / int(r[iX]) < 18]
    sm = sum(int(r[iN]) for r in lo); ex = sum(int(r[iX]) for r in lo); th = sum(int(r[iT]) for r in lo)
    print(f"\nInside the loop, instructions running at fewer than 18 lanes (leaf preparation, triangle filter, FP64 confirmation): "
          f"{100 * sm / tot:.1f} % of the samples, {100 * ex / totx:.1f} % of the warp instructions at {th / max(ex, 1):.1f} lanes.")
    step = max(int(r[iX]) for r in data[n0:n0 + 4])
    print(f"Node steps per launch (executions of the node fetch): {step / 1e6:.2f} M warp-steps.")
