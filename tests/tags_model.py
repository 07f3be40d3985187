"""Restatement of share/mergeTags/mergeTags.c (stranded mode) as a static link graph per overlap length: the fixed point
of the reference's pass-by-pass pairwise merging joins every PATH of links (mergeTags.c:74-205); representative = the
member with the lowest p-value, the rightmost on ties (:147-153 folded left to right). Test infrastructure: documents the
semantics dk_merge_tags implements and is itself checked against the unmodified binary (oracle/_ref/mergeTags)."""
import gzip
def assemble(a, k, min_k=15):
    # a: list of dict(seq, p, nb, rep)
    cyc = 0
    for i in range(k - 1, min_k - 1, -1):
        left, right = {}, {}
        def add(h, key, j):
            h[key] = j if key not in h else None
        for j, c in enumerate(a):
            if len(c["seq"]) <= i: continue
            add(left, c["seq"][:i], j); add(right, c["seq"][-i:], j)
        succ, pred = {}, {}
        for key, j in right.items():
            if j is None: continue
            l = left.get(key)
            if l is None or l == j: continue
            succ[j] = l; pred[l] = j
        used = set(); new = []
        for j in range(len(a)):
            if j in pred or j not in succ: continue
            cur = dict(a[j]); used.add(j); x = j
            while x in succ:
                y = succ[x]; used.add(y); r = a[y]
                cur = dict(seq=cur["seq"] + r["seq"][i:], nb=cur["nb"] + r["nb"], p=cur["p"] if cur["p"] < r["p"] else r["p"],
                           rep=cur["rep"] if cur["p"] < r["p"] else r["rep"])
                x = y
            new.append(cur)
        for j in range(len(a)):
            if j not in used:
                if j in succ and j in pred: cyc += 1
                new.append(a[j])
        a = new
    return a, cyc
def model(path, k=31):
    lines = gzip.open(path, "rt").read().split("\n")
    if lines and lines[-1] == "": lines.pop()
    hdr, body = lines[0], lines[1:]
    up, down = [], []
    for n, l in enumerate(body):
        f = l.split("\t")
        c = dict(seq=f[0], p=float(f[1]), nb=1, rep=n)
        (up if float(f[4]) > 0 else down).append(c)
    ua, c1 = assemble(up, k); da, c2 = assemble(down, k)
    out = ["nb_merged_kmers\tcontig\t" + hdr]
    for c in sorted(ua + da, key=lambda c: c["rep"]):
        out.append("%d\t%s\t%s" % (c["nb"], c["seq"], body[c["rep"]]))
    return "\n".join(out) + "\n", c1 + c2
