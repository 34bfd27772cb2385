"""Host-side GML front end (lexer, parser, evaluator).

Mirrors the public surface of /root/reference/internal/gml that callers of the
`render` boundary use: ``NewEvalState``, ``EvalState.ParseAndEval[File]``, the
``EvalState.Render`` hook, ``RenderArgs``, ``EvalSurfaceFn``, the scene-node and
light value types.  Out of scope for acceleration; present so
tests and benchmarks can start from the same GML text as the reference.
"""
from .evaluator import (BUILTINS, EXTENSION_BUILTINS, REFERENCE_BUILTINS, Cone, Cube, Cylinder, Difference, DirectionalLight, ErrArrayIndexOutOfBounds,
                        ErrEmptyStack, ErrTypeMismatch, ErrUnboundIdentifier, EvalState, EvalSurfaceFn, GMLError, Intersect,
                        Material, NewEvalState, Plane, PointLight, RenderArgs, SceneObject, Sphere, SpotLight, Union, VArray,
                        VClosure, VSurfaceFn, env_string, maybe_simplify_surface_fn, referenced_vars, value_string)
from .parser import GMLSyntaxError, IDMapping, Parser, format_float, token_list_string, token_string
from .lexer import Lexer, Token
