#!/usr/bin/env python
"""Operand-limited FP64 cost of a kernel's hottest loop, from its SASS (no GPU needed).

Model measured with tools/fp64_mix.cu and tools/fp64_reuse.cu on a B200: an FP64-pipe instruction (DADD, DMUL,
DFMA, DSETP) costs 2.0 scheduler cycles per warp, unless it reads THREE distinct register pairs from the register
file, then 2.96 (calibrated on six patterns of tools/fp64_reuse.cu, profiles/archive/r01_fp64_reuse_v9.md).  An operand that the previous instruction of the warp left in the operand reuse cache (`.reuse`
on the same register in the same slot) is not read from the register file.

    python tools/sass_fp64_cost.py gravidy_b200/libgravidy_b200.so k_force_warpILi2ELi4ELi1ELi3ELb0
    python tools/sass_fp64_cost.py gravidy_b200/libgravidy_b200.so k_force_warpILi2ELi1ELi4ELi3ELb0 4 --list
        (--list: the loop's SASS, every FP64 instruction marked  "3src" = three distinct register sources,
         "fed" = one of them comes from the operand reuse cache, "2.96" / "2.00" = modelled cycles)
"""
import re
import subprocess
import sys

FP64 = ("DADD", "DMUL", "DFMA", "DSETP")


def functions(path):
    out = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True, check=True).stdout
    name, body, res = None, [], {}
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            if name:
                res[name] = body
            name, body = m.group(1), []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m and name:
            body.append((int(m.group(1), 16), m.group(2).strip()))
    if name:
        res[name] = body
    return res


def hottest_loop(body):
    """the innermost loop (backward branch with no other loop inside) that holds the most FP64 instructions"""
    loops = []
    for addr, ins in body:
        m = re.search(r"BRA\S*\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)", ins)
        if m and int(m.group(1), 16) < addr:
            loops.append((int(m.group(1), 16), addr))
    best = None
    for tgt, addr in loops:
        if any((t2, a2) != (tgt, addr) and tgt <= t2 and a2 <= addr for t2, a2 in loops):
            continue
        seg = [(a, i) for a, i in body if tgt <= a <= addr]
        n = sum(1 for _, i in seg if re.sub(r"^@!?U?P\d\s+", "", i).startswith(FP64))
        if best is None or n > best[0]:
            best = (n, seg)
    return best[1] if best else body


def regs_of(ins):
    """(opcode, [source operand strings]) of an FP64 instruction"""
    ins = re.sub(r"^@!?U?P\d\s+", "", ins)
    op, rest = ins.split(None, 1)
    ops = [o.strip() for o in rest.split(",")]
    return op, ops[0], ops[1:]


def cost(seg, three=2.96, others_flush=True, marks=None):
    total, n64, n3, n3saved, prev_reuse = 0.0, 0, 0, 0, {}
    for addr, ins in seg:
        bare = re.sub(r"^@!?U?P\d\s+", "", ins)
        if not bare.startswith(FP64):
            # an instruction in between keeps the cache only if it does not touch the slots; be conservative
            if others_flush:
                prev_reuse = {}
            if marks is not None:
                marks.append((addr, ins, ""))
            continue
        op, dst, srcs = regs_of(ins)
        if op.startswith("DSETP"):
            srcs = [s for s in srcs if re.match(r"-?\|?R\d+", s)]
        n64 += 1
        distinct, cached, now_reuse = set(), set(), {}
        for slot, s in enumerate(srcs):
            m = re.match(r"[-|~]*R(\d+)(\.reuse)?", s)
            if not m:
                continue
            r = m.group(1)
            if prev_reuse.get(slot) == r:
                cached.add(r)
            if m.group(2):
                now_reuse[slot] = r
            distinct.add(r)
        read = distinct - cached
        if len(distinct) >= 3:
            n3 += 1
            if len(read) < 3:
                n3saved += 1
        total += three if len(read) >= 3 else 2.0
        if marks is not None:
            tagm = ("3src " if len(distinct) >= 3 else "     ") + ("fed  " if len(distinct) >= 3 and len(read) < 3 else "     ")
            marks.append((addr, ins, tagm + ("2.96" if len(read) >= 3 else "2.00")))
        prev_reuse = now_reuse
    return total, n64, n3, n3saved


def other_reg_readers(seg):
    """Instructions of other pipes that read a vector register (LDS / MUFU / integer with R operands): each takes one
    operand-read cycle away from the FP64 pipe (profiles/r02_force_lab.md; measured by removing them one by one).  Uniform
    datapath instructions (UIADD3, UISETP, LDS [UR + imm], BRA.U) and moves of RZ do not."""
    n = 0
    for _, ins in seg:
        bare = re.sub(r"^@!?U?P\d\s+", "", ins)
        if bare.startswith(FP64) or bare.startswith(("BRA", "NOP", "WARPSYNC", "BSYNC", "BSSY")):
            continue
        parts = bare.split(None, 1)
        if len(parts) < 2:
            continue
        ops = [o.strip() for o in parts[1].split(",")]
        srcs = ops[1:] if not parts[0].startswith(("ST", "RED", "ATOM")) else ops
        if any(re.search(r"\bR\d+", o) for o in srcs):
            n += 1
    return n


def main():
    path, pat = sys.argv[1], sys.argv[2]
    pairs = float(sys.argv[3]) if len(sys.argv) > 3 and not sys.argv[3].startswith("-") else 0.0
    listing = "--list" in sys.argv
    for name, body in functions(path).items():
        if pat not in name:
            continue
        seg = hottest_loop(body)
        total, n64, n3, n3saved = cost(seg)
        total_o, _, _, n3saved_o = cost(seg, others_flush=False)
        msg = (f"{name}: loop of {len(seg)} instr, {n64} FP64, {n3} with three register sources, "
               f"{n3saved} of them fed by the reuse cache; modelled {total:.1f} cycles per trip")
        if pairs:
            msg += f" = {total / pairs:.2f} per pair ({n64 / pairs:.1f} FP64 instr per pair)"
        msg += (f"; if LDS/MUFU/integer instructions in between do not flush the cache: {n3saved_o} fed, "
                f"{total_o / (pairs or 1):.2f}")
        nother = other_reg_readers(seg)
        msg += (f"; {nother} instructions of other pipes read a vector register (+1 cycle each): "
                f"{total + nother:.1f} cycles per trip" + (f" = {(total + nother) / pairs:.2f} per pair" if pairs else ""))
        print(msg)
        if listing:
            marks = []
            cost(seg, marks=marks)
            for addr, ins, tagm in marks:
                print(f"    /*{addr:05x}*/  {tagm:16s} {ins}")


if __name__ == "__main__":
    main()
