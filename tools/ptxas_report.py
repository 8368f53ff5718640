"""Developer helper: per-kernel registers / spills / shared memory from an `nvcc -Xptxas -v` log."""
import re, subprocess, sys
txt = open(sys.argv[1]).read()
only_spills = len(sys.argv) > 2 and sys.argv[2] == "spills"
for b in re.split(r"ptxas info\s+: Compiling entry function '", txt)[1:]:
    name = b.split("'")[0]
    m = re.search(r"Used (\d+) registers", b)
    sp = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", b)
    sm = re.search(r"(\d+) bytes smem", b)
    if only_spills and (not sp or sp.group(2) == "0"):
        continue
    dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    dem = dem.replace("void nb::force_kernel", "fk").replace("(nb::StepParams)", "").replace("(int)", "").replace("(bool)", "")
    print(dem[:110], "regs", m.group(1) if m else "?", "stack/spill", sp.groups() if sp else None, "smem", sm.group(1) if sm else 0)
