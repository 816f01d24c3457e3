"""Scratch: instruction mix of the sub-step loop of hb_wtile_kernel instantiations, from the SASS of the built objects
(no GPU needed).  The sub-step loop is taken to be the shortest backward loop that holds the whole RK4.   python tools/sass_loop_mix.py heat_b200/csrc/build/hb_wtile_2.o [substring of the mangled name ...]"""
import collections
import re
import subprocess
import sys

obj, filters = sys.argv[1], sys.argv[2:]
txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)[1:]
for f in funcs:
    name = f.split("\n", 1)[0].strip()
    if "hb_wtile_kernel" not in name or (filters and not any(x in name for x in filters)):
        continue
    ins = []   # (addr, opcode, text)
    for m in re.finditer(r"/\*([0-9a-f]{4,6})\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)([^;]*);", f):
        ins.append((int(m.group(1), 16), m.group(2), m.group(3)))
    addr_index = {a: i for i, (a, _, _) in enumerate(ins)}
    best = None
    for i, (a, op, rest) in enumerate(ins):
        if op.startswith("BRA"):
            t = re.search(r"0x([0-9a-f]+)", rest)
            if t and int(t.group(1), 16) < a and int(t.group(1), 16) in addr_index:
                j = addr_index[int(t.group(1), 16)]
                body = ins[j:i + 1]
                fp64 = sum(1 for _, o, _ in body if o.startswith(("DFMA", "DADD", "DMUL", "DSETP", "DMNMX")))
                # the sub-step loop: the SHORTEST backward loop that holds the whole RK4 (16 FP64 instructions per node slot)
                nl = int(re.search(r"hb_wtile_kernelILi\d+ELi(\d+)E", name).group(1))
                if fp64 >= 16 * nl and (best is None or len(body) < len(best[1])):
                    best = (fp64, body)
    if not best:
        continue
    hist = collections.Counter(o.split(".")[0] for _, o, _ in best[1])
    t = re.search(r"hb_wtile_kernelILi(\d+)ELi(\d+)ELi(\d+)ELb(\d)", name)
    total = len(best[1])
    groups = {"FP64": ("DFMA", "DADD", "DMUL", "DSETP", "DMNMX"), "select/predicate": ("FSEL", "SEL", "ISETP", "FSETP", "PLOP3", "LOP3"),
              "shared ld/st": ("LDS", "STS"), "local ld/st (spills)": ("LDL", "STL"), "shuffle": ("SHFL",), "MUFU/convert": ("MUFU", "F2F", "I2F", "F2I"),
              "move": ("MOV", "IMAD", "IADD3", "UMOV", "PRMT")}
    line = f"G={t.group(1)} NL={t.group(2)} MODE={t.group(3)}: {total} instructions in the sub-step loop; "
    line += "; ".join(f"{g} {sum(hist[o] for o in ops)}" for g, ops in groups.items())
    print(line)
    print("   per node slot:", round(total / int(t.group(2)), 1), "instructions,", round(sum(hist[o] for o in groups['FP64']) / int(t.group(2)), 1), "FP64")
