"""Joins two kernel_bench outputs (reference library, ours) into one table: per case the two times, the
ratio reference/ours (> 1: ours is faster) and whether the output checksums are identical."""
import json
import sys


def load(path, lib=None):
    rows, sums = {}, {}
    for ln in open(path):
        ln = ln.strip()
        if not ln.startswith("{"):
            continue
        try:
            d = json.loads(ln)
        except json.JSONDecodeError:
            continue
        if "case" not in d or (lib is not None and d.get("lib") != lib):
            continue
        (sums if "checksums" in d else rows)[d["case"]] = d
    return rows, sums


def main(ref_path, ours_path=None):
    """One file with both libraries (lib = "reference" / "ours", interleaved timing) or one file each."""
    ref, ref_s = load(ref_path, "reference" if ours_path is None else None)
    ours, ours_s = load(ours_path or ref_path, "ours" if ours_path is None else None)
    out = []
    print(f"{'case':46s} {'reference us':>13s} {'ours us':>9s} {'ref/ours':>9s} {'ours rate':>12s}  bytes equal")
    for case, o in ours.items():
        r = ref.get(case, {})
        same = None
        if case in ref_s and case in ours_s:
            same = {k: ref_s[case]["checksums"].get(k) == v for k, v in ours_s[case]["checksums"].items()}
        ru, ou = r.get("us"), o.get("us")
        ratio = (ru / ou) if ru and ou else None
        out.append({"case": case, "kernel": o.get("kernel"), "ref_us": ru, "ours_us": ou, "ref_over_ours": ratio,
                    "ours_rate": o.get("rate"), "unit": o.get("unit"), "bytes_equal": same,
                    "ref_error": r.get("error"), "ours_error": o.get("error")})
        eq = "-" if same is None else ("yes" if all(same.values()) else "NO: " + ",".join(k for k, v in same.items() if not v))
        print(f"{case:46s} {ru if ru else float('nan'):13.2f} {ou if ou else float('nan'):9.2f} "
              f"{ratio if ratio else float('nan'):9.3f} {o.get('rate', float('nan')):8.1f} {o.get('unit', ''):4s} {eq}"
              + (f"  [ref error: {r['error'][:90]}]" if r.get("error") else ""))
    return out


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[2].endswith(".jsonl"):
        res = main(sys.argv[1], sys.argv[2])
        out = sys.argv[3] if len(sys.argv) > 3 else None
    else:
        res = main(sys.argv[1])
        out = sys.argv[2] if len(sys.argv) > 2 else None
    if out:
        json.dump(res, open(out, "w"), indent=1)
