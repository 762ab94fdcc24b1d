"""List global loads that ptxas/nvcc placed BEFORE griddepcontrol.wait (SASS: ACQBULK) in each kernel of
libcbx_t3.so.  const __restrict__ loads are free to move above the wait's memory clobber; anything the
previous kernel produces must not appear in this list (weights / constants may)."""
import re, subprocess, sys
so = sys.argv[1] if len(sys.argv) > 1 else "chatterbox_vllm_b200/libcbx_t3.so"
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
fn, pre, seen_wait, out = None, [], False, {}
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn, pre, seen_wait = m.group(1), [], False
        out[fn] = (pre, [False])
        continue
    if fn is None:
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if not m:
        continue
    ins = m.group(2).strip()
    if "ACQBULK" in ins:
        seen_wait = True
        out[fn][1][0] = True
    elif not seen_wait and re.search(r"\b(LDG|LD\.E|UBLKCP|UTMALDG|LDGSTS)", ins):
        pre.append(f"{m.group(1)}: {ins}")
for fn, (pre, w) in out.items():
    name = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip()
    name = name.replace("(anonymous namespace)::", "").replace("void ", "").split("(")[0][:90]
    if w[0] and pre:
        print(f"{name}: {len(pre)} load(s) before the wait")
        for p in pre[:3]:
            print("    ", p)
    elif not w[0]:
        print(f"{name}: no griddepcontrol.wait")
    else:
        print(f"{name}: clean")
