"""Per-source-line view of an .ncu-rep captured with -lineinfo / --import-source on:

    python tests/devtools/ncu_lines.py gpurun_out/x.ncu-rep [top_n]

prints, per CUDA source line, the warp-instructions executed, the stall samples and the top stall
reasons (shares of the kernel totals), hottest lines first."""
import collections
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 60
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = None
    cur_file, cur = "?", None
    inst, samp, text = collections.Counter(), collections.Counter(), {}
    stalls = collections.defaultdict(collections.Counter)
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
            continue
        if r and r[0] == "Line No":
            hdr = r
            i_samp, i_inst = hdr.index("# Samples"), hdr.index("Instructions Executed")
            st_cols = [(i, h[6:]) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
            continue
        if hdr is None or len(r) < len(hdr):
            continue
        if r[0] != "":                    # a CUDA source line header: its own columns are the sums of the SASS below
            cur = (cur_file, int(r[0]))
            text[cur] = r[1].strip()
            try:
                inst[cur] += int(r[i_inst]); samp[cur] += int(r[i_samp])
            except ValueError:
                pass
            for i, name in st_cols:
                try:
                    stalls[cur][name] += int(r[i])
                except ValueError:
                    pass
    ti, ts = sum(inst.values()), sum(samp.values())
    print(f"total warp-instructions {ti}, samples {ts}")
    print(f"{'file:line':28s} {'inst%':>6s} {'samp%':>6s}  top stalls | source")
    for key, _ in sorted(samp.items(), key=lambda kv: -kv[1])[:top]:
        s3 = ", ".join(f"{n} {100 * v / max(samp[key], 1):.0f}%" for n, v in stalls[key].most_common(3))
        print(f"{key[0] + ':' + str(key[1]):28s} {100 * inst[key] / ti:6.2f} {100 * samp[key] / ts:6.2f}  {s3} | {text[key][:90]}")


if __name__ == "__main__":
    main()
