#!/usr/bin/env python
"""Annotates `objdump -dr` of one libm object file with the VALUES of the rodata constants each instruction
loads (resolving R_X86_64_PC32 relocations against .LC* / data symbols).  Study aid used while restating
glibc 2.39's x86-64 FMA libm variants for billiard_ai_b200/csrc/bcu_libm.cuh; not part of the product."""
import re
import struct
import subprocess
import sys


def sections(obj):
    """name -> bytes via objcopy-less parsing of `objdump -s`."""
    out = subprocess.check_output(["objdump", "-s", obj]).decode()
    secs, cur = {}, None
    for ln in out.splitlines():
        m = re.match(r"Contents of section (\S+):", ln)
        if m:
            cur = m.group(1); secs[cur] = {}
            continue
        m = re.match(r"^ ([0-9a-f]+) ((?:[0-9a-f]+ ?)+)", ln)
        if m and cur:
            off = int(m.group(1), 16)
            hexs = ln[len(m.group(1)) + 2:].split("  ")[0].replace(" ", "")
            # objdump prints up to 4 groups of 8 hex digits
            hexs = "".join(ln[1:].split()[1:5]) if True else hexs
            hexs = re.sub(r"[^0-9a-f]", "", "".join(ln.split("  ")[0].split()[1:]))
            secs[cur][off] = bytes.fromhex(hexs)
    flat = {}
    for name, chunks in secs.items():
        if not chunks:
            continue
        end = max(o + len(b) for o, b in chunks.items())
        buf = bytearray(end)
        for o, b in chunks.items():
            buf[o:o + len(b)] = b
        flat[name] = bytes(buf)
    return flat


def symbols(obj):
    out = subprocess.check_output(["objdump", "-t", obj]).decode()
    syms = {}
    for ln in out.splitlines():
        f = ln.split()
        if len(f) >= 4 and re.match(r"^[0-9a-f]{16}$", f[0]):
            name = f[-1]
            sec = f[-3] if len(f) >= 5 else None
            for tok in f:
                if tok.startswith(".rodata") or tok.startswith(".data") or tok == ".text.fma":
                    sec = tok
            syms[name] = (sec, int(f[0], 16))
    return syms


def main(obj, extra_objs):
    secs = sections(obj)
    syms = symbols(obj)
    ext = {}
    for eo in extra_objs:
        s2, y2 = sections(eo), symbols(eo)
        for name, (sec, off) in y2.items():
            if sec in s2:
                ext[name] = (s2[sec], off)
    out = subprocess.check_output(["objdump", "-dr", "-M", "intel", "--no-show-raw-insn", obj]).decode().splitlines()
    for i, ln in enumerate(out):
        m = re.search(r"R_X86_64_PC32\s+(\S+?)([+-]0x[0-9a-f]+)?$", ln)
        if m:
            name, add = m.group(1), int(m.group(2) or "0", 16)
            val = ""
            data = None
            if name in syms and syms[name][0] in secs:
                data, base = secs[syms[name][0]], syms[name][1]
            elif name in ext:
                data, base = ext[name]
            elif name in secs:
                data, base = secs[name], 0
            if data is not None:
                off = base + add + 4
                if 0 <= off and off + 8 <= len(data):
                    d = struct.unpack_from("<d", data, off)[0]
                    q = struct.unpack_from("<Q", data, off)[0]
                    val = f"   ; {name}+{add + 4:#x} = {d!r} ({q:#018x})"
                else:
                    val = f"   ; {name}+{add + 4:#x}"
            print(ln + val)
        else:
            print(ln)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2:])
