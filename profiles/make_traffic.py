"""profiles/traffic.json from an ncu launch list of the bench command:

    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
        -k regex:"warp_|rotate_|encode_|decode_" --csv --log-file gpurun_out/launches.csv \
        python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-fused --no-configs --no-sweep
    python profiles/make_traffic.py gpurun_out/launches.csv [cfg] [persons]

Per stage: dram__bytes_read.sum + dram__bytes_write.sum of one launch (mean over the captured launches of the
stage's kernels; the crop is the TMA kernel plus the cp.async kernel).  The file records a hash of the kernel
sources it was captured from; bench.py uses it for `roofline.traffic` only while that hash is the tree's."""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

STAGE_OF = (("warp_affine_u8_tma", "crop"), ("warp_affine_u8_tile", "crop"), ("warp_affine_u8_kernel", "crop"),
            ("encode_", "encode"), ("decode_", "decode"))


def main():
    path = sys.argv[1]
    cfg = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    persons = int(sys.argv[3]) if len(sys.argv) > 3 else 16384
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    ik, im, iv, iid = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("ID")
    per_launch = {}
    for r in rows[1:]:
        if r[im] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            per_launch.setdefault((r[iid], r[ik]), 0.0)
            unit = r[hdr.index("Metric Unit")]
            mul = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
            per_launch[(r[iid], r[ik])] += float(r[iv].replace(",", "")) * mul
    by_kernel = {}
    for (_, k), v in per_launch.items():
        by_kernel.setdefault(k.split("(")[0].split("<")[0].split("::")[-1].replace("void ", "").strip(), []).append(v)
    stage = {}
    for k, vs in by_kernel.items():
        for pat, st in STAGE_OF:
            if k.startswith(pat):
                stage[st] = stage.get(st, 0.0) + sum(vs) / len(vs)
                break
    out = {"note": "dram__bytes_read.sum + dram__bytes_write.sum per launch at %d persons, mean over the launches captured by "
                   "`ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none` on "
                   "`python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-fused --no-configs --no-sweep` (%s)" % (persons, os.path.basename(path)),
           "source_sha16": bench.kernel_source_sha16(), "persons": persons,
           "cfg%d" % cfg: {k: int(v) for k, v in sorted(stage.items())},
           "kernels": {k: int(sum(v) / len(v)) for k, v in sorted(by_kernel.items())}}
    json.dump(out, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
