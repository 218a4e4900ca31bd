"""Material parameter graphs: the PRL-level names of pine's Nodef / Node3f (src/pine/core/node.h:105-250,
node_context in node.cpp:29-132) and their flattening to the postfix programs of include/pine_b200.h
(pb2_node_instr). A graph is built with the reference's constructors and operators --

    Uber(lerp(Checkerboard(UV(), 0.95), [0.5, 0.7, 1.0], [0.01, 0.02, 0.03]), lerp(Checkerboard(UV(), 0.95), 0.0, 0.4))

(scenes/classic.pine:4-7) -- and `Node.program()` lists its instructions. NodeNoisef / NodeNoise3f and
NodeFunction (a PRL closure) have no device form and no constructor here.
"""
import numpy as np

OPS = {"end": 0, "const": 1, "position": 2, "normal": 3, "uv": 4, "+": 5, "-": 6, "*": 7, "/": 8, "^": 9,
       "neg": 10, "abs": 11, "sqr": 12, "sqrt": 13, "fract": 14, "component": 15, "vec3": 16, "checkerboard": 17,
       "image": 18, "imagef": 19}
FLOAT_FLAG = 0x100  # PB2_NODE_F: the instruction's value is a Nodef (all three components equal)


class Node:
    """One node of a graph. `is_float`: a Nodef (Mnode<float>), else a Node3f (Mnode<vec3>)."""

    def __init__(self, op, args=(), a=(0.0, 0.0, 0.0), is_float=False):
        self.op, self.args, self.a, self.is_float = op, tuple(args), tuple(float(np.float32(v)) for v in a), is_float

    # ---- operators registered in node_context (node.cpp:72-92) ----
    def _bin(self, op, other, swap=False):
        other = as_node(other)
        a, b = (other, self) if swap else (self, other)
        if a.is_float and b.is_float:
            return Node(op, (a, b), is_float=True)
        if op in ("+", "-") and a.is_float != b.is_float:
            raise TypeError(f"no `{op}` between a Nodef and a Node3f (node.cpp:72-92)")
        if op == "^" and a.is_float and not b.is_float:
            raise TypeError("no `^` with a Nodef base and a Node3f exponent (node.cpp:90)")
        return Node(op, (a, b), is_float=False)   # the Nodef operand becomes NodeToVec3(x): a broadcast

    def __add__(self, o): return self._bin("+", o)
    def __radd__(self, o): return self._bin("+", o, True)
    def __sub__(self, o): return self._bin("-", o)
    def __rsub__(self, o): return self._bin("-", o, True)
    def __mul__(self, o): return self._bin("*", o)
    def __rmul__(self, o): return self._bin("*", o, True)
    def __truediv__(self, o): return self._bin("/", o)
    def __rtruediv__(self, o): return self._bin("/", o, True)
    def __pow__(self, o): return self._bin("^", o)
    def __xor__(self, o): return self._bin("^", o)   # PRL spells pow as `^`
    def __neg__(self): return Node("neg", (self,), is_float=self.is_float)

    def __getitem__(self, k):   # ctx("[]") = NodeComponent (node.cpp:93)
        return Comp(self, k)

    def program(self, texture_index=None):
        """Postfix instruction list [(op | flag, a0, a1, a2)] WITHOUT the terminating end instruction."""
        out = []
        for x in self.args:
            out += x.program()
        code = OPS[self.op] | (FLOAT_FLAG if self.is_float else 0)
        if self.op == "vec3" and len(self.args) == 1:
            return out   # NodeToVec3(x): the identity on a stack whose floats are broadcast vec3s
        out.append((code,) + self.a)
        return out


def as_node(v):
    """int / float -> NodeConstant<float>, 3-sequence -> NodeConstant<vec3> (the ctor_variant lists, node.cpp:94-105)."""
    if isinstance(v, Node):
        return v
    if np.isscalar(v):
        f = float(np.float32(v))
        return Node("const", a=(f, f, f), is_float=True)
    v = [float(np.float32(c)) for c in v]
    if len(v) != 3:
        raise TypeError("a Node3f constant needs three components")
    return Node("const", a=v)


def is_graph(v):
    """True when a material parameter needs a program (anything but a plain constant)."""
    return isinstance(v, Node) and v.op != "const"


def constant_of(v, want_float):
    """The constant value of a parameter given as a number, a sequence or a constant Node; None for a graph."""
    n = as_node(v)
    if n.op != "const":
        return None
    return n.a[0] if want_float else n.a


def Position(): return Node("position")
def Normal(): return Node("normal")
def UV(): return Node("uv")


def Comp(x, k):
    x = as_node(x)
    if x.is_float:
        raise TypeError("Comp takes a Node3f")
    if int(k) not in (0, 1, 2):
        raise ValueError(f"NodeComponent's second parameter should be 0, 1, or 2, but get {k}")  # node.h:183-184
    return Node("component", (x,), a=(int(k), 0, 0), is_float=True)


def Vec3(x, y=None, z=None):
    if y is None:
        return Node("vec3", (_f(x),))
    return Node("vec3", (_f(x), _f(y), _f(z)))


def _f(x):
    x = as_node(x)
    if not x.is_float:
        raise TypeError("expected a Nodef")
    return x


def _v(x):
    x = as_node(x)
    return Vec3(x) if x.is_float else x


def Checkerboard(p, ratio=0.5):
    return Node("checkerboard", (_v(p),), a=(ratio, 0, 0), is_float=True)


def Texture(p, texture):
    """ctx.type<NodeImage>("Texture") (node.cpp:64): the texel of scene texture `texture` (index from
    Scene.add_texture) at fract(p.xy), nearest with wrap, gamma-decoded (node.cpp:20-23, image.cpp:7-18)."""
    return Node("image", (_v(p),), a=(int(texture) + 1, 0, 0))


def Texturef(p, texture):
    """NodeImagef (node.cpp:24-27): the same texel's first channel as a Nodef."""
    return Node("imagef", (_v(p),), a=(int(texture) + 1, 0, 0), is_float=True)


def _unary(op):
    def make(x):
        x = as_node(x)
        return Node(op, (x,), is_float=x.is_float)
    return make


abs_ = _unary("abs")
sqr = _unary("sqr")
sqrt = _unary("sqrt")
fract = _unary("fract")


def lerp(t, a, b):
    """ctx("lerp") (node.cpp:94-110): t * b + (1 - t) * a, built from the same binary nodes."""
    t, a, b = as_node(t), as_node(a), as_node(b)
    if t.is_float and a.is_float and b.is_float:
        return (t * b) + ((as_node(1.0) - t) * a)
    if a.is_float or b.is_float:
        raise TypeError("lerp(t, a, b): a and b must both be Nodef or both Node3f")
    if t.is_float:
        return (Vec3(t) * b) + (Vec3(as_node(1.0) - t) * a)
    return (t * b) + ((as_node((1.0, 1.0, 1.0)) - t) * a)
