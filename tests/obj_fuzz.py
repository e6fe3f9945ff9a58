"""Seeded generator of OBJ/MTL files for the live comparison of Scene::load with the reference's loader (tests/test_host_scene.py):
number spellings (plain, exponent, signs, leading/trailing dots, overflow/underflow, junk), polygons, negative and mixed indices,
groups/objects/usemtl switches (known and unknown names), comments, blank lines, tabs, trailing blanks, CRLF."""
import os

EXOTIC = [".5", "-.25", "5.", "+.5e1", "1e", "1e+", "-1.e2", "1.5e-", "0.1e+02", "12345678901234567890", "0.000000000000000000001234",
          "1.7976931348623157e308", "4.9e-324", "1e-400", "1e400", "00012.500", "-0", "+0.0", "3.4028235e38", "1.17549435e-38", "1.0000001",
          "0.30000000000000004", "123456789.123456789", "9007199254740993", "1e22", "1e23", "8.5e-5", "2.2250738585072011e-308", "1.e", "-", "+", "e5", "abc", "1,5", "1.2.3", "0x10", "1f", "1d5", "nan", "inf", "-inf"]
def num(r, level):
    if level >= 3 and r.random() < 0.25: return r.choice(EXOTIC)
    kind = r.randrange(8 if level > 0 else 3)
    v = r.uniform(-100, 100)
    if kind == 0: return "%d" % int(v)
    if kind == 1: return "%.3f" % v
    if kind == 2: return repr(v)
    if kind == 3: return "%.4e" % v
    if kind == 4: return ("+" if v >= 0 else "") + "%.2f" % v
    if kind == 5: return "%.5E" % (v * 1e-3)
    if kind == 6: return "%d." % int(v)
    return "%.6g" % (v * 1e3)

def gen(r, d, name, level):
    sep = lambda: r.choice([" ", " ", "  ", "\t"]) if level > 0 else " "
    nl = "\r\n" if (level > 1 and r.random() < 0.3) else "\n"
    mats = ["m%d" % i for i in range(r.randrange(0, 4))]
    lines = []
    if mats or r.random() < 0.3:
        lines.append("mtllib %s.mtl" % name)
        with open(os.path.join(d, name + ".mtl"), "w") as f:
            for m in mats:
                f.write("newmtl %s%s" % (m, nl))
                for key in r.sample(["Ka", "Kd", "Ks", "Ke", "Tf", "Kt"], r.randrange(1, 6)):
                    f.write("%s %s %s %s%s" % (key, num(r, 0), num(r, 0), num(r, 0), nl))
                for key in r.sample(["Ns", "Ni", "d", "Tr", "illum"], r.randrange(0, 5)):
                    f.write("%s %s%s" % (key, "%d" % r.randrange(0, 5) if key == "illum" else num(r, 0), nl))
                if r.random() < 0.3: f.write("map_Kd tex.png%s" % nl)
                if r.random() < 0.3: f.write("# comment%s%s" % (nl, nl))
    nv = r.randrange(3, 60)
    nn = r.randrange(0, 5); nt = r.randrange(0, 5)
    for i in range(nv):
        lines.append("v" + sep() + sep().join(num(r, level) for _ in range(3)))
        if level > 0 and r.random() < 0.05: lines.append("# c")
    for i in range(nn): lines.append("vn " + " ".join(num(r, 0) for _ in range(3)))
    for i in range(nt): lines.append("vt " + " ".join(num(r, 0) for _ in range(2)))
    nf = r.randrange(1, 40)
    for i in range(nf):
        if r.random() < 0.15: lines.append(r.choice(["g", "o"]) + " grp%d" % i)
        if r.random() < 0.2 and (mats or level > 0): lines.append("usemtl " + (r.choice(mats) if mats and r.random() < 0.8 else "nosuch"))
        k = r.choice([3, 3, 3, 4, 5, 6]) if level > 0 else 3
        idx = []
        for _ in range(k):
            vi = r.randrange(1, nv + 1)
            if level > 0 and r.random() < 0.2: vi = vi - nv - 1
            form = r.randrange(4)
            if form == 1 and nt: s = "%d/%d" % (vi, r.randrange(1, nt + 1))
            elif form == 2 and nn: s = "%d//%d" % (vi, r.randrange(1, nn + 1))
            elif form == 3 and nn and nt: s = "%d/%d/%d" % (vi, r.randrange(1, nt + 1), r.randrange(1, nn + 1))
            else: s = "%d" % vi
            idx.append(s)
        lines.append("f" + sep() + sep().join(idx) + (" " if level > 1 and r.random() < 0.1 else ""))
        if level > 1 and r.random() < 0.05: lines.append("")
    with open(os.path.join(d, name + ".obj"), "w", newline="") as f:
        f.write(nl.join(lines) + (nl if r.random() < 0.7 else ""))
    return os.path.join(d, name + ".obj")

