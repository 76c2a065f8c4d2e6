"""Condenses an ncu launch list (`ncu --metrics gpu__time_duration.sum --clock-control none --csv` of the bench command) into
  * a trimmed CSV (id, kernel, grid, block, ns) kept under profiles/, and
  * a JSON with the per-class shares of the steady-state forwards (every forward after the first, which also packs the
    q|k|v weights once) and a check that no `at::` (PyTorch) kernel runs between a forward's first patch-embedding launch
    and its unpatchify.
    python tools/launch_shares.py gpurun_out/launches.csv profiles/launches_r02_bench_c3.csv profiles/launch_shares_r02.json"""
import csv
import json
import re
import sys

CLASSES = [("attn_self", r"flash_attn_sa_kernel|flash_attn_kernel"), ("attn_cross", r"flash_attn_x\d_kernel"),
           ("gemm", r"gemm_bf16"), ("rows", r"ln_modulate|rmsnorm_rope"), ("patch_embed_unpatchify", r"patch_embed_kernel|patchify_kernel"),
           ("time_embed", r"sinusoid_kernel|gemv_bf16w|modulation_table")]


def short(n):
    n = re.sub(r"void |<unnamed>::|\(anonymous namespace\)::", "", n)
    return n.split("(")[0][:80]


def main(src, dst_csv, dst_json):
    with open(src) as f:
        lines = [l for l in f if l.startswith('"')]
    r = csv.reader(lines)
    hdr = next(r)
    ix = {h: i for i, h in enumerate(hdr)}
    rows = [(int(x[ix["ID"]]), short(x[ix["Kernel Name"]]), x[ix["Grid Size"]], x[ix["Block Size"]], int(float(x[ix["Metric Value"]])))
            for x in r]
    with open(dst_csv, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["id", "kernel", "grid", "block", "gpu__time_duration.sum [ns]"])
        w.writerows(rows)
    names = [x[1] for x in rows]
    starts = [i for i, n in enumerate(names) if n.startswith("patch_embed_kernel") and not names[i - 1].startswith("patch_embed_kernel")]
    ends = [i for i, n in enumerate(names) if n.startswith("unpatchify_kernel") and
            (i + 1 == len(names) or not names[i + 1].startswith("unpatchify_kernel"))]
    forwards = [(a, next(e for e in ends if e > a)) for a in starts if any(e > a for e in ends)]
    per, total, torch_inside = {}, 0.0, []
    steady = forwards[1:]
    for a, b in steady:
        torch_inside.append(sum(1 for i in range(a, b + 1) if names[i].startswith("at::")))
        for i in range(a, b + 1):
            cls = next((c for c, pat in CLASSES if re.search(pat, names[i])), "other")
            d = per.setdefault(cls, {"launches": 0, "ms": 0.0})
            d["launches"] += 1
            d["ms"] += rows[i][4] / 1e6
            total += rows[i][4] / 1e6
    for d in per.values():
        d["share"] = round(d["ms"] / total, 4)
        d["ms"] = round(d["ms"], 1)
    out = {"source": f"{dst_csv} (ncu --metrics gpu__time_duration.sum --clock-control none; serialised, cold cache, clocks not "
                     f"power-capped per kernel)",
           "launches_listed": len(rows), "forwards_found": len(forwards),
           "first_forward_launches": forwards[0][1] - forwards[0][0] + 1 if forwards else None,
           "steady_forwards": len(steady), "launches_per_steady_forward": [b - a + 1 for a, b in steady],
           "torch_kernels_between_first_patch_embed_and_unpatchify": torch_inside,
           "per_class": per, "total_ms": round(total, 1)}
    with open(dst_json, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(*sys.argv[1:4])
