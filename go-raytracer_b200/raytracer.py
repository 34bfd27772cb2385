"""Python mirror of the reference's root package API for the `render` path
(/root/reference/raytracer.go): ``ConvertRenderArgsToScene`` (:724), ``Render`` (:589),
``ParseAndRenderGML`` (:684), ``ParseAndRenderGMLFile`` (:692).  Same names, same argument
meaning, same error behaviour (lowering errors are raised; multiple / zero renders are errors
in ParseAndRenderGML*) - but the work is done by libgmlrt.so on the GPU.  No CPU path."""
from . import cuda, gml, lowering

_contexts = {}


def _context(devices, precision):
    key = (tuple(devices), precision)
    if key not in _contexts:
        _contexts[key] = cuda.Context(devices, precision)
    return _contexts[key]


class Scene:
    """What ConvertRenderArgsToScene returns: the lowered scene blob + where to render it."""

    def __init__(self, blob, width, height, devices=(0,), precision=cuda.PRECISION_F32):
        self.blob, self.WidthPx, self.HeightPx = blob, width, height
        self.devices, self.precision = devices, precision
        self.stats = None


def ConvertRenderArgsToScene(args, state=None, devices=(0,), precision=cuda.PRECISION_F32):
    """raytracer.go:724-754.  Raises lowering.LoweringError for unknown scene objects
    (the reference returns `unknown scene object type`, :825-826)."""
    return Scene(lowering.lower(args, state), args.Width, args.Height, devices, precision)


def Render(scene):
    """raytracer.go:589-682 -> uint8 array [H, W, 4] with the layout of image.RGBA.Pix."""
    img, scene.stats = _context(scene.devices, scene.precision).render(scene.blob)
    return img


class RenderError(Exception):
    pass


def _render_from_eval_state(state, run, devices, precision):
    images = {}

    def hook(st, args):
        images[args.File] = Render(ConvertRenderArgsToScene(args, st, devices, precision))

    state.Render = hook
    run()
    if len(images) > 1:
        raise RenderError("multiple images were rendered by the GML program")
    for img in images.values():
        return img
    raise RenderError("no image was rendered by the GML program")


def ParseAndRenderGML(program_text, devices=(0,), precision=cuda.PRECISION_F32):
    state = gml.NewEvalState()
    return _render_from_eval_state(state, lambda: state.ParseAndEval(program_text), devices, precision)


def ParseAndRenderGMLFile(path, devices=(0,), precision=cuda.PRECISION_F32):
    state = gml.NewEvalState()
    return _render_from_eval_state(state, lambda: state.ParseAndEvalFile(path), devices, precision)
