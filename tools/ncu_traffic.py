#!/usr/bin/env python3
"""DRAM traffic and warp-instruction counts per kernel from an ncu --set full report: dram__bytes_read.sum +
dram__bytes_write.sum and smsp__inst_executed.sum per launch, normalised per image (per pair for the per-pair kernels) so
bench.py can scale them to the launch it times (roofline.traffic, roofline.issue).
Usage: python tools/ncu_traffic.py gpurun_out/x.ncu-rep <images per launch> > profiles/rN/traffic.json"""
import csv
import json
import subprocess
import sys

TAGS = {"resize_kernel": "resize", "fast_score_kernel": "fast", "fast_score_win_kernel": "fast", "match_tc_kernel": "match_stereo", "fast_vote_kernel": "fast_cells", "octree_kernel": "octree",
        "blur_kernel": "blur", "describe_kernel": "describe", "describe_tma_kernel": "describe", "match_kernel": "match_stereo",
        "stereo_filter_kernel": "stereo_filter", "circular_kernel": "circular", "repack_kernel": "ingest"}


def main():
    rep, nimg = sys.argv[1], int(sys.argv[2])
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    ik, ir, iw, it = (hdr.index(k) for k in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"))
    ii = hdr.index("smsp__inst_executed.sum") if "smsp__inst_executed.sum" in hdr else None
    PER_PAIR = {"match_stereo", "match_temporal", "stereo_filter", "circular"}   # one problem per pair, not per image
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    out = {}
    seen_match = 0
    for r in rows[2:]:
        name = next((k for k in sorted(TAGS, key=len, reverse=True) if k + "(" in r[ik] or k + "<" in r[ik]), None)
        tag = TAGS.get(name)
        if tag is None:
            continue
        if tag == "match_stereo":
            seen_match += 1
            if seen_match % 2 == 0:
                tag = "match_temporal"
        b = float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]]
        e = out.setdefault(tag, {"dram_bytes_per_image": 0.0, "launches": 0, "us": 0.0})
        e["dram_bytes_per_image"] += b / nimg
        e["launches"] += 1
        if ii is not None:
            inst = float(r[ii].replace(",", ""))
            if tag in PER_PAIR:
                e["inst_per_pair"] = e.get("inst_per_pair", 0.0) + inst / (nimg / 2)
            else:
                e["inst_per_image"] = e.get("inst_per_image", 0.0) + inst / nimg
        e["us"] += float(r[it]) * (1e-3 if units[it] == "ns" else 1.0)   # report unit is ns or us
    json.dump({"source": rep.split("/")[-1], "images_per_launch": nimg,
               "note": "one step of the device-resident leg under ncu --set full; a kernel launched several times per step "
                       "(resize x7) is summed", "kernels": out}, sys.stdout, indent=1)


if __name__ == "__main__":
    main()
