#!/usr/bin/env python
"""ncu launch list (--csv, metrics gpu__time_duration.sum, dram__bytes_*.sum, smsp__inst_executed.sum) -> per-level JSON.

Sums the metrics per kernel name and over one level (init_kernel ... pack_records_kernel; one rank: no record merge).  Two levels are reported: the
one bench.py times (preceded by its L2 flush: the write-back of the flush buffer's dirty lines lands in this level's DRAM
write count) under "after_l2_flush", and the level before it (no flush in front: the path's own traffic) at top level.  bench.py reads `dram_bytes_read + dram_bytes_write` of the file it is pointed
at as roofline.traffic.   usage: ncu_level_json.py launches.csv "source note" > profiles/rNN_dram_per_level.json"""
import collections, csv, json, re, sys


def main(path, note):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))][1:]
    launches = collections.OrderedDict()
    for r in rows:
        name = re.sub(r"\(.*", "", r[4]).replace("unnamed>::", "")
        name = re.sub(r"_(dense|wide)$", "", name)   # the kernels for large parts: same work per item / job
        v = float(r[14].replace(",", ""))
        if r[13] == "ns":
            v /= 1e3                                                        # times in us
        launches.setdefault(int(r[0]), dict(kernel=name))[r[12]] = v
    seq = list(launches.values())
    flush = [i for i, l in enumerate(seq) if l["kernel"] == "flush_kernel"][0]
    ends = [i for i, l in enumerate(seq) if l["kernel"] == "pack_records_kernel"]
    starts = [i for i, l in enumerate(seq) if l["kernel"] == "init_kernel"]
    timed = seq[flush + 1:min(e for e in ends if e > flush) + 1]
    end = max(e for e in ends if e < flush)
    warm = seq[max(s for s in starts if s < end):end + 1]
    out = summarise(warm, note)
    out["after_l2_flush"] = summarise(timed, "the level bench.py times, right after flush_kernel")
    json.dump(out, sys.stdout, indent=1)
    print()


def summarise(level, note):
    per = collections.OrderedDict()
    for l in level:
        d = per.setdefault(l["kernel"], collections.defaultdict(float))
        d["launches"] += 1
        for k, v in l.items():
            if k != "kernel":
                d[k] += v
    tot = lambda m: sum(l.get(m, 0.0) for l in level)
    out = dict(source=note, launches=len(level), dram_bytes_read=tot("dram__bytes_read.sum"),
               dram_bytes_write=tot("dram__bytes_write.sum"), warp_instructions=tot("smsp__inst_executed.sum"),
               serialised_us=tot("gpu__time_duration.sum"),
               per_kernel={k: dict(v) for k, v in per.items()})
    return out


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
