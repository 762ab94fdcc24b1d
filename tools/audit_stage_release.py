"""Static audit of every shared-memory ring consumer in libcbx_t3.so: at the point where a warp RELEASES a stage
(SASS `SYNCS.ARRIVE.TRANS64.A1T0` = mbarrier.arrive without a transaction count), no LDS / LDSM of that warp
may still be in flight — the TMA engine refills the stage as soon as the barrier completes, and a shared-memory
read that has been issued but not performed would then see the next page (write-after-read across proxies).
ptxas is free to schedule the arrive ABOVE the instructions that consume the last load (it did, in the prefill
attention kernel: LDSM, SYNCS.ARRIVE, HMMA).  A release is accepted when, walking back from it to the previous
barrier wait / loop head, every shared load's destination registers are read by some instruction in between,
or a MEMBAR sits between the load and the arrive.

    python tools/audit_stage_release.py [libcbx_t3.so]     -> one line per kernel: "<kernel>: ok (n releases)" or "<kernel>: UNSAFE ..."
"""
import re
import subprocess
import sys

so = sys.argv[1] if len(sys.argv) > 1 else "chatterbox_vllm_b200/libcbx_t3.so"
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
kernels, cur = {}, None
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = kernels.setdefault(m.group(1), [])
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m and cur is not None:
        cur.append((m.group(1), m.group(2).strip()))


def regs(tok, width=1):
    m = re.match(r"R(\d+)", tok)
    return set() if not m else {int(m.group(1)) + i for i in range(width)}


def dest_regs(ins):
    """Destination register set of an LDS / LDSM."""
    m = re.search(r"\b(LDSM|LDS)(\S*)\s+R(\d+)", ins)
    if not m:
        return set()
    op, mods, r = m.group(1), m.group(2), int(m.group(3))
    n = 1
    if op == "LDSM":
        n = 4 if mods.endswith(".4") else (2 if mods.endswith(".2") else 1)
    elif ".128" in mods:
        n = 4
    elif ".64" in mods:
        n = 2
    return {r + i for i in range(n)}


def src_regs(ins):
    """Registers an instruction READS (everything after the first operand; wide reads over-approximated by 4)."""
    body = ins.split(None, 1)[1] if " " in ins else ""
    ops = body.split(",")
    out = set()
    for o in (ops[1:] if not ins.startswith(("ST", "@")) or True else ops):
        for m in re.finditer(r"R(\d+)", o):
            out |= {int(m.group(1)) + i for i in range(4)}
    if re.match(r"(@!?U?P\d+\s+)?(STS|STG|ST\.|ATOMS|RED)", ins):          # stores read their "first" data operand too
        for m in re.finditer(r"R(\d+)", body):
            out |= {int(m.group(1)) + i for i in range(4)}
    return out


bad = 0
for fn, ins in kernels.items():
    rel = [i for i, (_, s) in enumerate(ins) if "SYNCS.ARRIVE" in s and ".A1T0" in s]
    if not rel:
        continue
    name = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip()
    name = name.replace("(anonymous namespace)::", "").replace("void ", "").split("(")[0][:90]
    problems = []
    for i in rel:
        pending = []                                         # (addr, text, dest regs) of loads not yet seen consumed
        j = i - 1
        consumed_after = set()                               # registers read between the position and the arrive
        fenced = False
        while j >= 0 and i - j < 400:
            a, s = ins[j]
            if "MEMBAR" in s:
                fenced = True
            if "SYNCS.PHASECHK" in s or ("SYNCS.ARRIVE" in s and ".A1T0" in s):
                break                                        # previous wait / release: the window of this stage use
            d = dest_regs(s)
            if d and not fenced and not (d & consumed_after):
                problems.append(f"{a}: {s}  (still in flight at the release {ins[i][0]})")
            consumed_after |= src_regs(s)
            j -= 1
    if problems:
        bad += 1
        print(f"{name}: UNSAFE release of a stage")
        for p in problems[:4]:
            print("    ", p)
    else:
        print(f"{name}: ok ({len(rel)} release(s))")
sys.exit(1 if bad else 0)
