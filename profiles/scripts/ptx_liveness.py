"""Registers that are live around the back edge of the outermost loop of a PTX kernel.

Usage: python profiles/scripts/ptx_liveness.py file.ptx kernel_name_substring [label]

Why: the persistent codeblock kernel wraps 15 000 lines of straight-line code in a loop. A virtual register that is only
written under a condition (and read under the same condition later) is, for the register allocator, live on the path around
the loop that skips the write -- i.e. through every row body. This prints the virtual registers that are live-in at the loop
header (default: the target of the last backward branch of the kernel) together with the first line that defines and the
first line that reads each of them, so the source can be fixed (define the value on every path).

A plain iterative backward data-flow analysis over basic blocks; PTX syntax handled: predicated instructions (a predicated
write does not kill), vector operands {a, b}, setp's p|q destinations, st / bar / bra / red without destinations.
"""
import re
import sys
from collections import defaultdict

REG = re.compile(r"%(?:rd|rs|r|p|f|fd|h|hh|b)\d+")
NO_DEST = ("st.", "bar.", "barrier.", "bra", "ret", "red.", "membar", "fence", "exit", "trap", "prefetch", "tcgen05.st",
           "tcgen05.wait", "tcgen05.dealloc", "tcgen05.relinquish", "tcgen05.fence", "cp.async", "mbarrier.init", "brx.idx",
           "call", "nanosleep", "griddepcontrol", "tcgen05.alloc", "atom.shared.max", "pmevent")


def parse(path, kname):
    lines = open(path).read().split("\n")
    start = next(i for i, l in enumerate(lines) if ".entry" in l and kname in l)
    depth, body, began = 0, [], False
    for i in range(start, len(lines)):
        l = lines[i]
        depth += l.count("{") - l.count("}") if not l.strip().startswith("//") else 0
        if "{" in l:
            began = True
        body.append((i + 1, l))
        if began and depth == 0:
            break
    return body


def split_instr(body):
    """-> list of (lineno, label or None, text) with multi-line asm joined on ';'."""
    out, cur, cur_line = [], "", None
    for no, l in body:
        t = l.split("//")[0].strip()
        if not t or t.startswith(".") and not t.startswith(".reg") and ":" not in t:
            continue
        if t.startswith(".reg") or t.startswith(".local") or t.startswith(".shared") or t.startswith(".param"):
            continue
        if t in ("{", "}"):
            continue
        m = re.match(r"^(\$?[A-Za-z_][\w$]*):$", t)
        if m:
            out.append((no, m.group(1), None))
            continue
        # inline-asm scopes: "{ .reg .b16 tmp; mov.b32 {%rs1, tmp}, %r2; }" -- statement by statement, scope braces dropped
        for piece in re.findall(r"[^;]*;|[^;]+$", t):
            piece = piece.strip()
            while piece.startswith("}") or (piece.startswith("{") and not re.match(r"^\{[^}]*\}\s*,", piece)):
                piece = piece[1:].strip()
            if not piece or piece == ";" or piece.startswith(".reg"):
                continue
            if cur_line is None:
                cur_line = no
            cur += " " + piece
            if piece.endswith(";"):
                out.append((cur_line, None, cur.strip().rstrip(";")))
                cur, cur_line = "", None
    return out


def defs_uses(text):
    pred = None
    t = text
    m = re.match(r"^@(!?)(%p\d+)\s+(.*)$", t)
    if m:
        pred, t = m.group(2), m.group(3)
    op = t.split()[0]
    rest = t[len(op):]
    regs_all = REG.findall(rest)
    defs, uses = [], []
    if any(op.startswith(x) for x in NO_DEST):
        uses = regs_all
    else:
        # destination = first operand (may be {a,b,..} or p|q)
        first = rest.strip()
        if first.startswith("{"):
            dpart = first[:first.index("}") + 1]
            rem = first[len(dpart):]
        else:
            dpart = first.split(",")[0]
            rem = first[len(dpart):]
        if "[" in dpart:  # e.g. an instruction whose first operand is an address: no register destination
            uses = regs_all
        else:
            defs = REG.findall(dpart)
            uses = REG.findall(rem)
    if pred:
        uses = uses + [pred]
    return defs, uses, pred is not None, op


def main():
    path, kname = sys.argv[1], sys.argv[2]
    ins = split_instr(parse(path, kname))
    # basic blocks
    blocks, cur = [], {"label": "entry", "ins": []}
    for no, label, text in ins:
        if label is not None:
            if cur["ins"] or cur["label"] == "entry":
                blocks.append(cur)
            cur = {"label": label, "ins": []}
            continue
        cur["ins"].append((no, text))
        op = re.sub(r"^@!?%p\d+\s+", "", text).split()[0]
        if op.startswith("bra") or op.startswith("ret") or op.startswith("exit") or op.startswith("brx"):
            blocks.append(cur)
            cur = {"label": None, "ins": []}
    if cur["ins"]:
        blocks.append(cur)
    idx = {b["label"]: i for i, b in enumerate(blocks) if b["label"]}
    # brx.idx targets: declared in .branchtargets lines of the original text (collect all labels there)
    raw = open(path).read()
    succ = defaultdict(list)
    for i, b in enumerate(blocks):
        fall = True
        if b["ins"]:
            no, text = b["ins"][-1]
            t = re.sub(r"^@!?%p\d+\s+", "", text)
            op = t.split()[0]
            if op.startswith("bra"):
                tgt = t.split()[-1]
                if tgt in idx:
                    succ[i].append(idx[tgt])
                fall = text.startswith("@")
            elif op.startswith("brx"):
                name = t.split(",")[-1].strip()
                m = re.search(re.escape(name) + r":\s*\.branchtargets([^;]*);", raw)
                for tgt in re.findall(r"\$?[A-Za-z_][\w$]*", m.group(1)) if m else []:
                    if tgt in idx:
                        succ[i].append(idx[tgt])
                fall = False
            elif op.startswith("ret") or op.startswith("exit"):
                fall = text.startswith("@")
        if fall and i + 1 < len(blocks):
            succ[i].append(i + 1)
    # per-block use/def
    use, kill = [], []
    for b in blocks:
        u, k = set(), set()
        for no, text in b["ins"]:
            d, us, predicated, _ = defs_uses(text)
            for r in us:
                if r not in k:
                    u.add(r)
            if not predicated:
                k.update(d)
            else:
                for r in d:  # predicated write: old value may survive
                    if r not in k:
                        u.add(r)
        use.append(u)
        kill.append(k)
    live_in = [set() for _ in blocks]
    changed = True
    while changed:
        changed = False
        for i in range(len(blocks) - 1, -1, -1):
            out = set()
            for s in succ[i]:
                out |= live_in[s]
            new = use[i] | (out - kill[i])
            if new != live_in[i]:
                live_in[i] = new
                changed = True
    # loop headers: targets of backward branches. Without a label argument: list them (widest loops first) and stop.
    if len(sys.argv) > 3:
        hdr = idx[sys.argv[3]]
    else:
        loops = {}
        for i, b in enumerate(blocks):
            for s in succ[i]:
                if s <= i and b["ins"]:
                    first = blocks[s]["ins"][0][0] if blocks[s]["ins"] else 0
                    span = b["ins"][-1][0] - first
                    if s not in loops or span > loops[s][0]:
                        loops[s] = (span, first, b["ins"][-1][0])
        print("loops (header label, first line, last line of the back-edge block, registers live-in at the header):")
        for s, (span, first, last) in sorted(loops.items(), key=lambda kv: -kv[1][0])[:12]:
            print("  %-16s %6d .. %6d   %d" % (blocks[s]["label"], first, last, len(live_in[s])))
        return
    first_def, first_use = {}, {}
    for b in blocks:
        for no, text in b["ins"]:
            d, us, _, _ = defs_uses(text)
            for r in d:
                first_def.setdefault(r, (no, text))
            for r in us:
                first_use.setdefault(r, (no, text))
    regs = sorted(live_in[hdr], key=lambda r: (r.rstrip("0123456789"), int(re.search(r"\d+$", r).group())))
    n32 = sum(1 for r in regs if r.startswith("%r") and not r.startswith("%rd") and not r.startswith("%rs"))
    n64 = sum(1 for r in regs if r.startswith("%rd"))
    print("%d virtual registers live-in at the header (%d 32-bit, %d 64-bit, rest predicates / 16-bit)" % (len(regs), n32, n64))
    for r in regs:
        d = first_def.get(r, (0, "?"))
        u = first_use.get(r, (0, "?"))
        print("%-8s def line %-6d %-60s | first use line %-6d %s" % (r, d[0], d[1][:60], u[0], u[1][:60]))


if __name__ == "__main__":
    main()
