"""Markdown tables for DESIGN.md / BASELINE.md from bench.py records.

    python profiles/make_tables.py profiles/r02_bench_v4.json profiles/r02_bench_ref_v4.json \
           profiles/r02_bench_n2.json profiles/r02_bench_n4.json profiles/r02_bench_n8.json
"""
import json
import sys


def load(p):
    return json.loads(open(p).read().strip().splitlines()[-1])


def main():
    one, ref = load(sys.argv[1]), load(sys.argv[2])
    multi = [load(p) for p in sys.argv[3:]]
    B = one["config"]["persons_per_gpu_per_step"]
    rows = [dict(cfg=one["config"]["cfg"], ms_per_step=one["ms_per_step"], value=one["value"], kernels=one["kernels"],
                 encode_path_frac=one["paths"]["encode_path"]["frac"], decode_path_frac=one["paths"]["decode_path"]["frac"])] + one["configs"]
    rows.sort(key=lambda r: r["cfg"])
    print("### per config, one B200, %d persons per step\n" % B)
    print("| cfg | crop ms (frac) | encode ms (frac) | decode ms (frac) | encode path frac | decode path frac | step ms | M persons/s |")
    print("|---|---|---|---|---|---|---|---|")
    for r in rows:
        k = r["kernels"]
        print("| %d | %.2f (%.2f) | %.2f (%.2f) | %.2f (%.2f) | %.2f | %.2f | %.2f | %.2f |" % (
            r["cfg"], k["crop"]["ms"], k["crop"]["frac"], k["encode"]["ms"], k["encode"]["frac"], k["decode"]["ms"], k["decode"]["frac"],
            r["encode_path_frac"], r["decode_path_frac"], r["ms_per_step"], r["value"] / 1e6))
    e = one["e2e"]
    print("\ne2e (host buffers): %.1f k persons/s (%.1f k streaming); H2D %.0f B, D2H %.0f B per person; link ceiling %.1f k (H2D %.1f + D2H %.1f GB/s)" % (
        e["value"] / 1e3, e["streaming_value"] / 1e3, e["h2d_bytes_per_step"] / e["persons_per_step_per_gpu"], e["d2h_bytes_per_step"] / e["persons_per_step_per_gpu"],
        e["link"]["ceiling_persons_per_s"] / 1e3, e["link"]["h2d_gbs"], e["link"]["d2h_gbs"]))
    c = one.get("cpu_baseline", {})
    print("CPU arm (same run): %.0f persons/s on %s cores; --impl reference: %.0f persons/s; stage us/person/core %s" % (
        c.get("value", 0), c.get("cores"), ref["value"], ref["cpu_baseline"]["stage_us_per_person_per_core"]))
    nx = one.get("next", {})
    print("next rows:", {k: (v.get("ms"), v.get("frac", v.get("achieved_gbs"))) for k, v in nx.items() if isinstance(v, dict)})
    print("\n### scaling (weak, %d persons per GPU per step)\n" % B)
    print("| GPUs | M persons/s | x of 1 GPU | step ms | gather ms | e2e k persons/s | link sum H2D + D2H GB/s |")
    print("|---|---|---|---|---|---|---|")
    for d in [one] + multi:
        print("| %d | %.2f | %.2f | %.3f | %.3f | %.1f | %.0f + %.0f |" % (
            d["n_gpus"], d["value"] / 1e6, d["value"] / one["value"], d["ms_per_step"], d["gather_ms"], d["e2e"]["value"] / 1e3,
            d["e2e"]["link"]["h2d_gbs"], d["e2e"]["link"]["d2h_gbs"]))
    print("\n### cfg-5 sweep, strong scaling: ms (M persons/s)\n")
    hdr = [one] + multi
    print("| persons | " + " | ".join("N=%d" % d["n_gpus"] for d in hdr) + " | %d-GPU efficiency |" % hdr[-1]["n_gpus"])
    print("|---|" + "---|" * (len(hdr) + 1))
    for i, p in enumerate(one["sweep"]["points"]):
        cells = []
        for d in hdr:
            q = d["sweep"]["points"][i]
            cells.append("%.3f (%.2f)" % (q["ms"], q["persons_per_s"] / 1e6))
        last = hdr[-1]["sweep"]["points"][i]
        eff = p["ms"] / last["ms"] / hdr[-1]["n_gpus"]
        print("| %d | " % p["persons"] + " | ".join(cells) + " | %.2f |" % eff)
    print("\nclock records: headline", one["clocks"], "; sweep reasons", sorted({r for d in hdr for q in d["sweep"]["points"] for r in (q["clocks"] or {}).get("reasons", [])}))


def baseline_rows():
    """The rows of BASELINE.md section 6 (stage rows per config at one GPU, whole-step rows at 1/2/4/8 GPUs)."""
    one, ref = load(sys.argv[1]), load(sys.argv[2])
    multi = [load(p) for p in sys.argv[3:]]
    peak = one["roofline"]["peak"]
    rows = [dict(cfg=one["config"]["cfg"], value=one["value"], kernels=one["kernels"])] + one["configs"]
    rows.sort(key=lambda r: r["cfg"])
    cpu = ref["cpu_baseline"]
    print("\n### BASELINE.md section 6\n")
    print("| cfg | stage | GPUs | persons/s | achieved GB/s | %% of 8 TB/s | %% of measured peak (%.0f GB/s) | CPU persons/s/core | CPU persons/s/box (cores) | parity |" % peak)
    print("|---|---|---|---|---|---|---|---|---|---|")
    B = one["config"]["persons_per_gpu_per_step"]
    for r in rows:
        for st in ("crop", "encode", "decode"):
            k = r["kernels"][st]
            print("| %d | %s | 1 | %.2f M | %.0f | %.0f %% | %.0f %% | | | green |" % (r["cfg"], st, B / k["ms"] / 1e3, k["achieved_gbs"], k["achieved_gbs"] / 80.0, 100 * k["frac"]))
        print("| %d | whole step | 1 | %.2f M | | | | | | green |" % (r["cfg"], r["value"] / 1e6))
    for d in multi:
        print("| %d | whole step | %d | %.2f M | | | | | | green |" % (d["config"]["cfg"], d["n_gpus"], d["value"] / 1e6))
    print("| %d | whole step, host buffers (e2e) | 1 | %.1f k | | | | %.0f | %.0f (%d) | green |" % (
        one["config"]["cfg"], one["e2e"]["value"] / 1e3, cpu["per_core"], ref["value"], cpu["cores"]))
    for d in multi:
        print("| %d | whole step, host buffers (e2e) | %d | %.1f k | | | | | | green |" % (d["config"]["cfg"], d["n_gpus"], d["e2e"]["value"] / 1e3))


if __name__ == "__main__":
    main()
    baseline_rows()
