"""Scene lowering: *RenderArgs -> one flat, pointer-free scene blob.

This is the host half of what ``ConvertRenderArgsToScene`` does in the reference
(/root/reference/raytracer.go:724-830): per leaf ObjectToWorld = TransformMat (or
identity) and WorldToObject = Mat4.Inverse (vec.go:319-365).  The reference repeats
the conversion for each of its 8 worker goroutines and keeps Go pointers to closures;
here it runs once, surface closures are compiled to bytecode (bytecode.py) and the
result is the byte buffer described in include/gmlrt.h.  Union flattening, cube face
planes and normal matrices are derived from these bits on the library side.
"""
import struct

from . import prim
from .bytecode import Pools, SurfaceCompileError, compile_surface
from .gml import evaluator as ev

MAGIC, VERSION = 0x424C4D47, 1
HEADER_BYTES, NODE_BYTES, LIGHT_BYTES, SURFACE_BYTES = 256, 208, 96, 96

NODE_KIND = {ev.Union: 0, ev.Intersect: 1, ev.Difference: 2, ev.Sphere: 16, ev.Cube: 17, ev.Cylinder: 18, ev.Cone: 19,
             ev.Plane: 20}


class LoweringError(Exception):
    pass


def _rows(m):
    return [m[r][c] for r in range(3) for c in range(4)]


def lower(args, state=None):
    """RenderArgs -> bytes.  ``state`` is accepted for signature parity with
    ConvertRenderArgsToScene(args, state); closures carry their own environment."""
    pools = Pools()
    surfaces, surface_index = [], {}
    nodes = []

    def surface_of(fn):
        key = id(fn.material) if fn.material is not None else id(fn.closure)
        if key in surface_index:
            return surface_index[key]
        if fn.material is not None:
            rec = (0, 0, 0, 0, fn.material.as_tuple())
        else:
            try:
                res = compile_surface(fn.closure, pools)
            except SurfaceCompileError as e:
                raise LoweringError(f"surface function: {e}") from e
            if res[0] == "const":
                rec = (0, 0, 0, 0, res[1])
            else:
                rec = (1, res[1], res[2], res[3], res[4])
        surface_index[key] = len(surfaces)
        surfaces.append(rec)
        return surface_index[key]

    def add(root_obj):
        """Iterative (union chains of thousands of nodes are normal): node i is created when first
        seen, its operand indices are patched in once the operands have been numbered."""
        todo = [(root_obj, None, 0)]           # (object, parent node index, which operand)
        root_idx = None
        while todo:
            obj, parent, slot = todo.pop()
            kind = NODE_KIND.get(type(obj))
            if kind is None:
                raise LoweringError(f"unknown scene object type {type(obj).__name__}")
            idx = len(nodes)
            if parent is None:
                root_idx = idx
            else:
                nodes[parent][1 + slot] = idx
            if kind < 16:
                if isinstance(obj, ev.Union):
                    if len(obj.objects) != 2:
                        raise LoweringError("union nodes are binary")
                    a, b = obj.objects
                else:
                    a, b = obj.a, obj.b
                nodes.append([kind, 0, 0, 0, [0.0] * 12, [0.0] * 12])
                todo.append((b, idx, 1))
                todo.append((a, idx, 0))
            else:
                o2w = obj.transform_mat if obj.transform_mat is not None else prim.identity()
                w2o = prim.inverse(o2w) if obj.transform_mat is not None else prim.identity()
                if w2o is None:
                    raise LoweringError("singular transform (the reference dereferences nil here, vec.go:343-347)")
                nodes.append([kind, 0, 0, surface_of(obj.surface_fn), _rows(o2w), _rows(w2o)])
        return root_idx

    root = add(args.Scene)

    lights = []
    for l in args.Lights:
        if type(l) is ev.PointLight:
            lights.append((0, tuple(l.position), tuple(l.color), (0.0, 0.0, 0.0), 0.0, 0.0))
        elif type(l) is ev.DirectionalLight:
            lights.append((1, tuple(l.direction), tuple(l.color), (0.0, 0.0, 0.0), 0.0, 0.0))
        elif type(l) is ev.SpotLight:
            lights.append((2, tuple(l.position), tuple(l.color), tuple(l.at), l.cutoff, l.exp))
        else:
            raise LoweringError(f"expected lights array to contain *PointLight, got {type(l).__name__}")

    def align8(n):
        return (n + 7) & ~7

    off_nodes = HEADER_BYTES
    off_lights = off_nodes + NODE_BYTES * len(nodes)
    off_surfaces = off_lights + LIGHT_BYTES * len(lights)
    off_fpool = off_surfaces + SURFACE_BYTES * len(surfaces)
    off_itab = off_fpool + 8 * len(pools.fpool)
    off_code = align8(off_itab + 4 * len(pools.itab))
    total = align8(off_code + 4 * len(pools.code))

    out = bytearray(total)
    amb = tuple(args.AmbientLight)
    struct.pack_into("<IIQiiiId3d3d3dIIIIIIII6Q", out, 0, MAGIC, VERSION, total, args.Width, args.Height, args.Depth, 0,
                     float(args.Fov), *amb, *tuple(args.BgColorStart), *tuple(args.BgColorEnd),
                     len(nodes), root, len(lights), len(surfaces), len(pools.fpool), len(pools.itab), len(pools.code), 0,
                     off_nodes, off_lights, off_surfaces, off_fpool, off_itab, off_code)
    for i, (kind, a, b, s, o2w, w2o) in enumerate(nodes):
        struct.pack_into("<IIII12d12d", out, off_nodes + NODE_BYTES * i, kind, a, b, s, *o2w, *w2o)
    for i, (kind, pos, col, at, cutoff, exp) in enumerate(lights):
        struct.pack_into("<II3d3d3ddd", out, off_lights + LIGHT_BYTES * i, kind, 0, *pos, *col, *at, cutoff, exp)
    for i, (kind, coff, clen, nregs, mat) in enumerate(surfaces):
        struct.pack_into("<IIII10d", out, off_surfaces + SURFACE_BYTES * i, kind, coff, clen, nregs, *mat)
    struct.pack_into(f"<{len(pools.fpool)}d", out, off_fpool, *pools.fpool)
    struct.pack_into(f"<{len(pools.itab)}i", out, off_itab, *pools.itab)
    struct.pack_into(f"<{len(pools.code)}I", out, off_code, *pools.code)
    return bytes(out)


def describe(blob):
    """Decode a blob back into plain dicts (tests, debugging)."""
    h = struct.unpack_from("<IIQiiiId3d3d3dIIIIIIII6Q", blob, 0)
    keys = ("magic version total_bytes width height depth flags fov ax ay az s0 s1 s2 e0 e1 e2 n_nodes root n_lights "
            "n_surfaces n_fpool n_itab n_code reserved0 off_nodes off_lights off_surfaces off_fpool off_itab off_code").split()
    hd = dict(zip(keys, h))
    nodes = []
    for i in range(hd["n_nodes"]):
        r = struct.unpack_from("<IIII12d12d", blob, hd["off_nodes"] + NODE_BYTES * i)
        nodes.append(dict(kind=r[0], a=r[1], b=r[2], surface=r[3], o2w=r[4:16], w2o=r[16:28]))
    lights = []
    for i in range(hd["n_lights"]):
        r = struct.unpack_from("<II3d3d3ddd", blob, hd["off_lights"] + LIGHT_BYTES * i)
        lights.append(dict(kind=r[0], pos=r[2:5], color=r[5:8], at=r[8:11], cutoff=r[11], exponent=r[12]))
    surfaces = []
    for i in range(hd["n_surfaces"]):
        r = struct.unpack_from("<IIII10d", blob, hd["off_surfaces"] + SURFACE_BYTES * i)
        surfaces.append(dict(kind=r[0], code_off=r[1], code_len=r[2], n_regs=r[3], material=r[4:14]))
    fpool = struct.unpack_from(f"<{hd['n_fpool']}d", blob, hd["off_fpool"])
    itab = struct.unpack_from(f"<{hd['n_itab']}i", blob, hd["off_itab"])
    code = struct.unpack_from(f"<{hd['n_code']}I", blob, hd["off_code"])
    return dict(header=hd, nodes=nodes, lights=lights, surfaces=surfaces, fpool=fpool, itab=itab, code=code)
