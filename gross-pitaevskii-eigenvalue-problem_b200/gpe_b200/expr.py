"""Straight-line scalar programs over the moment vector(s) -- the host side of `gpe::scalar_program`.

The loss terms of the reference end in scalar arithmetic on reduced quantities (mu = <u,Hu>/<u,u>, mean r^2 with that mu
inside, 1/(mu^2 + 1e-6), the Riesz energy ...: NBroot cell 6:117-125, S/gross_pitaevskii_2D.py:192-211).  `losses.py`
writes that arithmetic once, with ordinary python operators; evaluated on `losses.Moments` it runs as float64 torch
scalars (one small kernel per operator, plus the autograd graph behind them), evaluated on `SymMoments` it only RECORDS
the operations.  `evaluate` compiles the record into an opcode table (cached on the device per structure and constants,
so nothing is uploaded inside a CUDA-graph capture) and one kernel returns every requested term AND its Jacobian with
respect to the moments; the backward of the terms is a single [T] x [T, n_in] product.
"""
import torch

from ._lib import MOM_GRAD2, MOM_NN2, MOM_SYM2, NMOM, F_ONE, mom

CONST, INPUT, ADD, SUB, MUL, DIV, NEG, SQRT = range(8)
MAX_NODES = 1024


class Expr:
    """A recorded scalar: node of a straight-line program (no value of its own)."""
    __slots__ = ("op", "a", "b", "c")

    def __init__(self, op, a=None, b=None, c=0.0):
        self.op, self.a, self.b, self.c = op, a, b, c

    @staticmethod
    def lift(v):
        if isinstance(v, Expr):
            return v
        if isinstance(v, torch.Tensor):
            raise TypeError("gpe_b200.expr: tensors cannot enter a recorded expression (only python numbers and moments)")
        return Expr(CONST, c=float(v))

    def __add__(self, o): return Expr(ADD, self, Expr.lift(o))
    def __radd__(self, o): return Expr(ADD, Expr.lift(o), self)
    def __sub__(self, o): return Expr(SUB, self, Expr.lift(o))
    def __rsub__(self, o): return Expr(SUB, Expr.lift(o), self)
    def __mul__(self, o): return Expr(MUL, self, Expr.lift(o))
    def __rmul__(self, o): return Expr(MUL, Expr.lift(o), self)
    def __truediv__(self, o): return Expr(DIV, self, Expr.lift(o))
    def __rtruediv__(self, o): return Expr(DIV, Expr.lift(o), self)
    def __neg__(self): return Expr(NEG, self)

    def __pow__(self, p):
        if p == 0.5:
            return Expr(SQRT, self)
        if int(p) != p or p < 1 or p > 8:
            raise ValueError("recorded expressions support integer powers 1..8 and 0.5")
        out = self
        for _ in range(int(p) - 1):
            out = Expr(MUL, out, self)
        return out

    def sqrt(self): return Expr(SQRT, self)


class SymMoments:
    """`losses.Moments` for recording: the same accessors, returning Expr nodes that read input vector `slot`."""

    def __init__(self, slot=0):
        self.base = slot * NMOM

    def _in(self, k):
        return Expr(INPUT, c=float(self.base + k))

    def P(self, i, j): return self._in(mom(i, j))
    @property
    def n(self): return self._in(mom(F_ONE, F_ONE))
    def mean(self, i, j=F_ONE): return self.P(i, j) / self.n
    @property
    def grad2(self): return self._in(MOM_GRAD2)
    @property
    def nn2(self): return self._in(MOM_NN2)
    @property
    def sym2(self): return self._in(MOM_SYM2)

    def quad(self, w):
        items = list(w.items())
        total = 0.0
        for a, (i, wi) in enumerate(items):
            for (j, wj) in items[a:]:
                total = total + (1.0 if i == j else 2.0) * wi * wj * self.P(i, j)
        return total

    def dot_u(self, w):
        from ._lib import F_U
        return sum(wi * self.P(F_U, i) for i, wi in w.items())


def _compile(outs):
    """Topological opcode table of the expressions `outs` (common subexpressions by identity)."""
    index, ops, consts = {}, [], []

    def visit(e):
        # iterative post-order: recursion depth would follow the length of a sum
        stack = [(e, False)]
        while stack:
            node, done = stack.pop()
            if id(node) in index:
                continue
            if not done:
                stack.append((node, True))
                for child in (node.a, node.b):
                    if child is not None and id(child) not in index:
                        stack.append((child, False))
            else:
                if node.op == CONST:
                    ops.append((CONST, 0, 0)); consts.append(node.c)
                elif node.op == INPUT:
                    ops.append((INPUT, int(node.c), 0)); consts.append(0.0)
                else:
                    ops.append((node.op, index[id(node.a)], index[id(node.b)] if node.b is not None else 0)); consts.append(0.0)
                index[id(node)] = len(ops) - 1
    out_idx = []
    for e in outs:
        e = Expr.lift(e)
        visit(e)
        out_idx.append(index[id(e)])
    if len(ops) > MAX_NODES:
        raise RuntimeError(f"gpe_b200.expr: {len(ops)} nodes exceed the kernel's {MAX_NODES}")
    return tuple(ops), tuple(consts), tuple(out_idx)


_programs = {}     # (key, device) -> (ops, consts, outs) device tensors


class _Program(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inputs, ops_t, consts_t, outs_t):
        vals, jac = torch.ops.gpe.scalar_program(ops_t, consts_t, inputs.contiguous(), outs_t)
        ctx.save_for_backward(jac)
        return vals

    @staticmethod
    def backward(ctx, g):
        (jac,) = ctx.saved_tensors
        return g.reshape(1, -1).mm(jac).reshape(-1), None, None, None


def evaluate(moment_vectors, build, key=None):
    """Terms `build(*SymMoments)` (a sequence of recorded scalars) evaluated on the float64 moment vectors (CUDA tensors of
    NMOM entries, one per SymMoments argument): a tuple of 0-dim float64 tensors, differentiable with respect to the
    vectors.  `key` (hashable; must determine the program: family and every constant in it) lets the compiled table be
    reused without re-recording."""
    dev = moment_vectors[0].device
    ck = (key, dev) if key is not None else None
    prog = _programs.get(ck) if ck is not None else None
    if prog is None:
        table = _compile(list(build(*[SymMoments(s) for s in range(len(moment_vectors))])))
        prog = _programs.get((table, dev))
        if prog is None:
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError("gpe_b200.expr: a loss program was first met inside a CUDA-graph capture; run the step once "
                                   "eagerly (GraphedStep's warm-up does) so its table is uploaded beforehand")
            ops, consts, outs = table
            prog = (torch.tensor(ops, dtype=torch.int32, device=dev).reshape(-1, 3).contiguous(),
                    torch.tensor(consts, dtype=torch.float64, device=dev), torch.tensor(outs, dtype=torch.int32, device=dev))
            _programs[(table, dev)] = prog
        if ck is not None:
            _programs[ck] = prog
    inputs = moment_vectors[0] if len(moment_vectors) == 1 else torch.cat(list(moment_vectors))
    vals = _Program.apply(inputs, *prog)
    return vals.unbind(0)
