"""profiles/sass_summary.txt: per-kernel SASS mnemonic counts of libtcb200.so (cuobjdump -sass)."""
import re
import subprocess
import sys

COLS = ["IMMA", "LDSM", "IDP", "UBLKCP", "UBLKPF", "SYNCS", "BAR", "PRMT", "LDG", "STG", "LDS", "STS", "LDL", "STL", "ATOMS", "ATOMG", "SHFL"]
out = subprocess.run(["cuobjdump", "-sass", "triton_client_b200/libtcb200.so"], capture_output=True, text=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
kernels, cur = {}, None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = demangle(m.group(1))
        cur = re.sub(r"\(.*", "", cur).replace("void ", "").replace("tcb::", "").replace("(int)", "").replace("(bool)", "")
        kernels[cur] = {c: 0 for c in COLS}
        kernels[cur]["instrs"] = 0
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        kernels[cur]["instrs"] += 1
        base = op.split(".")[0]
        for c in COLS:
            if base == c or (c == "ATOMS" and base in ("ATOMS", "ATOMS.CAS")) or (c == "ATOMG" and base in ("ATOMG", "ATOM", "RED")) \
                    or (c == "UBLKPF" and base == "UBLKPF"):
                kernels[cur][c] += 1
                break
all_ops = set(re.findall(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", out))
with open("profiles/sass_summary.txt", "w") as f:
    f.write("SASS instruction counts per kernel of triton_client_b200/libtcb200.so (cuobjdump -sass, sm_100a only; scratch/sass_summary.py).\n")
    f.write("IMMA = mma.sync int8 tensor path, LDSM = ldmatrix, IDP = dp4a, UBLKCP = cp.async.bulk (1-D TMA), UBLKPF = its L2 prefetch form, SYNCS = mbarrier ops.\n")
    tc = sorted(o for o in all_ops if o.startswith(("UTC", "UTMA", "LDTM", "STTM", "HMMA", "QMMA")))
    f.write("Tensor-map TMA / tcgen05 / HMMA mnemonics present in the library: %s\n\n" % (", ".join(tc) if tc else "none"))
    f.write("%-64s %8s" % ("kernel", "instrs") + "".join("%8s" % c for c in COLS) + "\n")
    for k in sorted(kernels):
        f.write("%-64s %8d" % (k[:64], kernels[k]["instrs"]) + "".join("%8d" % kernels[k][c] for c in COLS) + "\n")
print(len(kernels), "kernels;", "prefetch-like ops:", sorted(o for o in all_ops if "PF" in o or "PREF" in o))
