"""Straight-line DFT codelet generator (a small genfft).

Builds the real-arithmetic expression DAG of a length-R DFT from radix-2/3/4/5
butterflies composed with Good-Thomas (coprime factors, no twiddles) or
Cooley-Tukey (twiddles folded in as constants), with
  * hash-consing (common sub-expressions are emitted once),
  * sign tracking (negations are folded into add/sub),
  * constant folding of multiplications by 0, +-1 (and +-i for complex values),
  * zero-input pruning (real-input codelets, zero-padded frames),
  * dead-code elimination (only the requested outputs are emitted).
The emitter writes CUDA/C++ `__host__ __device__` inline functions whose a*b+c
patterns the compiler contracts into FFMA.

Used by tools/gen_kernels.py; the generated header is committed under
expressive_speech_translation_b200/csrc/generated/.
"""
import math
from fractions import Fraction


class Dag:
    """Real-valued expression DAG with hash-consing.  A value is (sign, node_id) or None (= 0)."""

    def __init__(self):
        self.nodes = []  # tuples: ('in', name) | ('add', a, b) | ('sub', a, b) | ('mul', c, a)
        self.memo = {}

    def _mk(self, key):
        i = self.memo.get(key)
        if i is None:
            i = len(self.nodes)
            self.nodes.append(key)
            self.memo[key] = i
        return i

    def inp(self, name):
        return (1, self._mk(("in", name)))

    def neg(self, x):
        return None if x is None else (-x[0], x[1])

    def add(self, x, y):
        if x is None:
            return y
        if y is None:
            return x
        (sx, ix), (sy, iy) = x, y
        if ix == iy:
            if sx == sy:
                return self.mul(2.0, x)
            return None
        if sx == sy:
            a, b = (ix, iy) if ix < iy else (iy, ix)
            return (sx, self._mk(("add", a, b)))
        # x - y  or  y - x
        if sx > 0:
            return (1, self._mk(("sub", ix, iy)))
        return (1, self._mk(("sub", iy, ix)))

    def sub(self, x, y):
        return self.add(x, self.neg(y))

    def mul(self, c, x):
        if x is None or c == 0.0:
            return None
        if abs(abs(c) - 1.0) < 1e-15:
            return x if c > 0 else self.neg(x)
        s = x[0] * (1 if c > 0 else -1)
        node = self.nodes[x[1]]
        if node[0] == "mul":  # fold constant chains
            c2 = abs(c) * node[1]
            if abs(c2 - 1.0) < 1e-15:
                return (s, node[2])
            return (s, self._mk(("mul", c2, node[2])))
        return (s, self._mk(("mul", abs(c), x[1])))


class C:
    """complex value over a Dag: (re, im) of Dag values."""

    __slots__ = ("re", "im")

    def __init__(self, re, im):
        self.re = re
        self.im = im


def cadd(d, a, b):
    return C(d.add(a.re, b.re), d.add(a.im, b.im))


def csub(d, a, b):
    return C(d.sub(a.re, b.re), d.sub(a.im, b.im))


def cmul_i(d, a, sign):
    """multiply by +i (sign=+1) or -i (sign=-1)."""
    if sign > 0:
        return C(d.neg(a.im), a.re)
    return C(a.im, d.neg(a.re))


def cscale(d, c, a):
    return C(d.mul(c, a.re), d.mul(c, a.im))


def _snap(v):
    for t in (0.0, 1.0, -1.0, 0.5, -0.5):
        if abs(v - t) < 1e-15:
            return t
    return v


def twiddle(num, den):
    """exp(-2*pi*i*num/den) with exact zeros/ones."""
    fr = Fraction(num, den) % 1
    ang = -2.0 * math.pi * float(fr)
    # exact octants
    e = fr * 8
    if e.denominator == 1:
        r = math.sqrt(0.5)
        tab = [(1, 0), (r, -r), (0, -1), (-r, -r), (-1, 0), (-r, r), (0, 1), (r, r)]
        return tab[int(e)]
    return (_snap(math.cos(ang)), _snap(math.sin(ang)))


def cmul_const(d, w, a):
    wr, wi = w
    if wi == 0.0:
        return cscale(d, wr, a)
    if wr == 0.0:
        return cscale(d, abs(wi), cmul_i(d, a, 1 if wi > 0 else -1))
    re = d.sub(d.mul(wr, a.re), d.mul(wi, a.im))
    im = d.add(d.mul(wr, a.im), d.mul(wi, a.re))
    return C(re, im)


# ---------------------------------------------------------------- small DFTs
def dft2(d, x):
    return [cadd(d, x[0], x[1]), csub(d, x[0], x[1])]


def dft3(d, x):
    c = -0.5
    s = math.sin(2.0 * math.pi / 3.0)
    t1 = cadd(d, x[1], x[2])
    t2 = csub(d, x[1], x[2])
    y0 = cadd(d, x[0], t1)
    a = cadd(d, x[0], cscale(d, c, t1))
    b = cscale(d, s, t2)  # y1 = a - i*b ; y2 = a + i*b
    mib = cmul_i(d, b, -1)
    return [y0, cadd(d, a, mib), csub(d, a, mib)]


def dft4(d, x):
    a = cadd(d, x[0], x[2])
    b = csub(d, x[0], x[2])
    c = cadd(d, x[1], x[3])
    e = csub(d, x[1], x[3])
    mie = cmul_i(d, e, -1)
    return [cadd(d, a, c), cadd(d, b, mie), csub(d, a, c), csub(d, b, mie)]


def dft5(d, x):
    c1 = math.cos(2.0 * math.pi / 5.0)
    c2 = math.cos(4.0 * math.pi / 5.0)
    s1 = math.sin(2.0 * math.pi / 5.0)
    s2 = math.sin(4.0 * math.pi / 5.0)
    t1 = cadd(d, x[1], x[4])
    t2 = cadd(d, x[2], x[3])
    t3 = csub(d, x[1], x[4])
    t4 = csub(d, x[2], x[3])
    y0 = cadd(d, cadd(d, x[0], t1), t2)
    # (x0 + c*t) + c'*t' : every add has one single-use product operand -> two FFMA per component
    a1 = cadd(d, cadd(d, x[0], cscale(d, c1, t1)), cscale(d, c2, t2))
    a2 = cadd(d, cadd(d, x[0], cscale(d, c2, t1)), cscale(d, c1, t2))
    b1 = cadd(d, cscale(d, s1, t3), cscale(d, s2, t4))
    b2 = csub(d, cscale(d, s2, t3), cscale(d, s1, t4))
    mib1 = cmul_i(d, b1, -1)
    mib2 = cmul_i(d, b2, -1)
    return [y0, cadd(d, a1, mib1), cadd(d, a2, mib2), csub(d, a2, mib2), csub(d, a1, mib1)]


BASE = {2: dft2, 3: dft3, 4: dft4, 5: dft5}

# preferred factorizations  R -> (R1, R2): coprime => Good-Thomas, else Cooley-Tukey
FACTOR = {6: (2, 3), 8: (2, 4), 10: (2, 5), 12: (3, 4), 15: (3, 5), 16: (4, 4), 20: (4, 5),
          24: (3, 8), 25: (5, 5), 30: (5, 6), 32: (4, 8), 40: (5, 8), 60: (4, 15), 64: (8, 8)}


def dft(d, x):
    """forward DFT of the list of complex Dag values x (natural order in, natural order out)."""
    n = len(x)
    if n == 1:
        return [x[0]]
    if n in BASE:
        return BASE[n](d, x)
    r1, r2 = FACTOR[n]
    if math.gcd(r1, r2) == 1:
        # Good-Thomas: n = (r2*n1 + r1*n2) mod N ; k = (k1*r2*inv(r2,r1) + k2*r1*inv(r1,r2)) mod N
        inner = []
        for n2 in range(r2):
            inner.append(dft(d, [x[(r2 * n1 + r1 * n2) % n] for n1 in range(r1)]))  # over n1 -> k1
        e1 = r2 * pow(r2, -1, r1)
        e2 = r1 * pow(r1, -1, r2)
        out = [None] * n
        for k1 in range(r1):
            col = dft(d, [inner[n2][k1] for n2 in range(r2)])  # over n2 -> k2
            for k2 in range(r2):
                out[(k1 * e1 + k2 * e2) % n] = col[k2]
        return out
    # Cooley-Tukey: n = r2*n1 + n2 ; k = k1 + r1*k2
    inner = []
    for n2 in range(r2):
        a = dft(d, [x[r2 * n1 + n2] for n1 in range(r1)])
        inner.append([cmul_const(d, twiddle(n2 * k1, n), a[k1]) for k1 in range(r1)])
    out = [None] * n
    for k1 in range(r1):
        col = dft(d, [inner[n2][k1] for n2 in range(r2)])
        for k2 in range(r2):
            out[k1 + r1 * k2] = col[k2]
    return out


# ---------------------------------------------------------------- evaluation / emission
def evaluate(d, outs, env):
    """numerically evaluate Dag values `outs` (list of (sign,id)|None) given env {name: float}."""
    val = [0.0] * len(d.nodes)
    for i, nd in enumerate(d.nodes):
        if nd[0] == "in":
            val[i] = env[nd[1]]
        elif nd[0] == "add":
            val[i] = val[nd[1]] + val[nd[2]]
        elif nd[0] == "sub":
            val[i] = val[nd[1]] - val[nd[2]]
        else:
            val[i] = nd[1] * val[nd[2]]
    return [0.0 if o is None else o[0] * val[o[1]] for o in outs]


def count_ops(d, outs):
    live = _live(d, outs)
    cnt = {"add": 0, "sub": 0, "mul": 0}
    for i in live:
        k = d.nodes[i][0]
        if k in cnt:
            cnt[k] += 1
    return cnt


def _live(d, outs):
    live = set()
    stack = [o[1] for o in outs if o is not None]
    while stack:
        i = stack.pop()
        if i in live:
            continue
        live.add(i)
        nd = d.nodes[i]
        if nd[0] in ("add", "sub"):
            stack.extend((nd[1], nd[2]))
        elif nd[0] == "mul":
            stack.append(nd[2])
    return sorted(live)


def _flit(c):
    s = repr(float(c))
    if "e" not in s and "." not in s and "inf" not in s:
        s += ".0"
    return s + "f"


def emit_body(d, outs, in_expr, out_stmt, indent="    ", vtype="V"):
    """in_expr: name -> C expression; out_stmt: (index, expr_string) -> statement string.
    outs: list of Dag values.  Returns C statements over the value type `vtype` (float or the packed
    two-frame type fe2) written with the fe_add / fe_sub / fe_mulc / fe_fmac / fe_fmsc / fe_fnmac / fe_neg
    helpers of fe_common.h.  A product with exactly one consumer that is an add/sub is contracted into
    that consumer as an explicit fused multiply-add (the packed intrinsics are never contracted by nvcc)."""
    live = _live(d, outs)
    uses = {i: 0 for i in live}
    for i in live:
        nd = d.nodes[i]
        if nd[0] in ("add", "sub"):
            uses[nd[1]] += 1
            uses[nd[2]] += 1
        elif nd[0] == "mul":
            uses[nd[2]] += 1
    for o in outs:
        if o is not None:
            uses[o[1]] += 1

    def fusable(j):
        return d.nodes[j][0] == "mul" and uses[j] == 1

    fused = set()
    plan = {}
    for i in live:
        nd = d.nodes[i]
        if nd[0] == "add":
            a, b = nd[1], nd[2]
            if fusable(b) and b not in fused:
                plan[i] = ("fmac", b, a)
                fused.add(b)
            elif fusable(a) and a not in fused:
                plan[i] = ("fmac", a, b)
                fused.add(a)
        elif nd[0] == "sub":
            a, b = nd[1], nd[2]
            if fusable(b) and b not in fused:
                plan[i] = ("fnmac", b, a)      # a - c*x
                fused.add(b)
            elif fusable(a) and a not in fused:
                plan[i] = ("fmsc", a, b)       # c*x - b
                fused.add(a)
    lines = []
    name = {}
    for i in live:
        nd = d.nodes[i]
        if nd[0] == "in":
            name[i] = in_expr(nd[1])
            continue
        if i in fused:
            continue
        v = f"t{i}"
        name[i] = v
        if i in plan:
            op, m, other = plan[i]
            mn = d.nodes[m]
            lines.append(f"{indent}const {vtype} {v} = fe_{op}({_flit(mn[1])}, {name[mn[2]]}, {name[other]});")
        elif nd[0] == "add":
            lines.append(f"{indent}const {vtype} {v} = fe_add({name[nd[1]]}, {name[nd[2]]});")
        elif nd[0] == "sub":
            lines.append(f"{indent}const {vtype} {v} = fe_sub({name[nd[1]]}, {name[nd[2]]});")
        else:
            lines.append(f"{indent}const {vtype} {v} = fe_mulc({_flit(nd[1])}, {name[nd[2]]});")
    for k, o in enumerate(outs):
        if o is None:
            e = f"fe_zero<{vtype}>()"
        else:
            e = name[o[1]] if o[0] > 0 else f"fe_neg({name[o[1]]})"
        st = out_stmt(k, e)
        if st:
            lines.append(indent + st)
    return "\n".join(lines)
