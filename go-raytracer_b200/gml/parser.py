"""GML token groups (AST) and the recursive-descent parser.

Mirrors /root/reference/internal/gml/expr.go:10-128 (node kinds) and
parser.go:43-226 (grammar: TokenList ::= TokenGroup*, TokenGroup ::= Token |
``{ TokenList }`` | ``[ TokenList ]``).  Host-side only.
"""
from . import lexer as lx


class GMLSyntaxError(Exception):
    pass


class Pos:
    __slots__ = ("line", "col")

    def __init__(self, line=0, col=0):
        self.line, self.col = line, col

    def prefix(self):
        return f"{self.line}:{self.col}: " if self.line else ""


class Node:
    __slots__ = ("pos",)


class Identifier(Node):
    __slots__ = ("name", "id")

    def __init__(self, name, id_, pos):
        self.name, self.id, self.pos = name, id_, pos


class Binder(Node):
    __slots__ = ("name", "id")

    def __init__(self, name, id_, pos):
        self.name, self.id, self.pos = name, id_, pos


class IntLiteral(Node):
    __slots__ = ("value",)

    def __init__(self, value, pos):
        self.value, self.pos = value, pos


class FloatLiteral(Node):
    __slots__ = ("value",)

    def __init__(self, value, pos):
        self.value, self.pos = value, pos


class BoolLiteral(Node):
    __slots__ = ("value",)

    def __init__(self, value, pos):
        self.value, self.pos = value, pos


class StringLiteral(Node):
    __slots__ = ("value",)

    def __init__(self, value, pos):
        self.value, self.pos = value, pos


class Function(Node):
    __slots__ = ("body",)

    def __init__(self, body, pos):
        self.body, self.pos = body, pos


class Array(Node):
    __slots__ = ("elements",)

    def __init__(self, elements, pos):
        self.elements, self.pos = elements, pos


def format_float(f):
    """expr.go:130-139: strconv.FormatFloat(f, 'g', -1, 64), plus a trailing ".0"
    when the result would otherwise look like an integer."""
    f = float(f)
    if f != f:
        return "NaN"
    if f in (float("inf"), float("-inf")):
        return "+Inf" if f > 0 else "-Inf"
    # shortest round-trip digits and decimal exponent from repr()
    mant, _, exp = repr(abs(f)).partition("e")
    whole, _, fracpart = mant.partition(".")
    digits = (whole + fracpart).lstrip("0")
    dp = len(whole) + (int(exp) if exp else 0)   # decimal point position, fixed up below
    if not digits:
        digits, dp = "0", 1
    else:
        lead = len(whole + fracpart) - len((whole + fracpart).lstrip("0"))
        dp -= lead
        digits = digits.rstrip("0") or "0"
    sign = "-" if (f < 0 or (f == 0 and str(f)[0] == "-")) else ""
    e10 = dp - 1
    if e10 < -4 or e10 >= 6:   # 'g' with shortest digits: exponent form outside [1e-4, 1e6)
        m = digits[0] + ("." + digits[1:] if len(digits) > 1 else "")
        s = f"{sign}{m}e{'+' if e10 >= 0 else '-'}{abs(e10):02d}"
    elif dp <= 0:
        s = sign + "0." + "0" * (-dp) + digits
    elif len(digits) <= dp:
        s = sign + digits + "0" * (dp - len(digits))
    else:
        s = sign + digits[:dp] + "." + digits[dp:]
    if any(ch in s for ch in ".eE"):
        return s
    return s + ".0"


def token_string(node):
    """expr.go:141-160 TokenGroupDebugString."""
    if isinstance(node, IntLiteral):
        return str(node.value)
    if isinstance(node, FloatLiteral):
        return go_float(node.value)
    if isinstance(node, BoolLiteral):
        return "true" if node.value else "false"
    if isinstance(node, StringLiteral):
        return '"' + node.value.replace("\\", "\\\\").replace('"', '\\"').replace("\n", "\\n").replace("\t", "\\t") + '"'
    if isinstance(node, Identifier):
        return node.name
    if isinstance(node, Binder):
        return "/" + node.name
    if isinstance(node, Function):
        return "{ " + token_list_string(node.body) + " }"
    if isinstance(node, Array):
        return "[ " + token_list_string(node.elements) + " ]"
    raise TypeError(f"unknown token group: {node!r}")


go_float = format_float


def token_list_string(nodes):
    return " ".join(token_string(n) for n in nodes)


class IDMapping:
    """environment.go:93-127: interned symbol ids shared across scopes."""

    def __init__(self):
        self.name_to_id, self.id_to_name, self.max_id = {}, {}, 0

    def get_or_create(self, name):
        i = self.name_to_id.get(name)
        if i is None:
            self.max_id += 1
            i = self.max_id
            self.name_to_id[name] = i
            self.id_to_name[i] = name
        return i

    def clone(self):
        m = IDMapping()
        m.name_to_id, m.id_to_name, m.max_id = dict(self.name_to_id), dict(self.id_to_name), self.max_id
        return m


_STARTS_GROUP = {lx.LBRACKET, lx.LCURLY, lx.IDENT, lx.INT, lx.FLOAT, lx.STRING, lx.BINDER, lx.BOOLEAN}


class Parser:
    def __init__(self, lexer, id_mapping=None):
        self.lexer = lexer
        self.ids = id_mapping if id_mapping is not None else IDMapping()
        self.cur = None

    @classmethod
    def from_text(cls, text, id_mapping=None):
        return cls(lx.Lexer(text), id_mapping)

    @classmethod
    def from_file(cls, path, id_mapping=None):
        return cls(lx.Lexer.from_file(path), id_mapping)

    def parse(self):
        self.cur = self.lexer.next_token()
        body = self._list()
        t = self.cur
        if t.kind == lx.ERROR:
            raise GMLSyntaxError(f"{t.line}:{t.col}: {t.text}")
        if t.kind != lx.EOF:
            raise GMLSyntaxError(f"{t.line}:{t.col}: unexpected token: {t.kind}, expected end of input")
        return body

    def _take(self):
        t = self.cur
        self.cur = self.lexer.next_token()
        return t

    def _expect(self, kind):
        t = self.cur
        if t.kind == lx.ERROR:
            raise GMLSyntaxError(f"{t.line}:{t.col}: {t.text}")
        if t.kind != kind:
            raise GMLSyntaxError(f"{t.line}:{t.col}: expected {kind}, got {t.kind}")
        self._take()

    def _list(self):
        out = []
        while self.cur.kind in _STARTS_GROUP:
            out.append(self._group())
        return out

    def _group(self):
        t = self.cur
        pos = Pos(t.line, t.col)
        k = t.kind
        if k == lx.LBRACKET:
            self._expect(lx.LBRACKET)
            body = self._list()
            self._expect(lx.RBRACKET)
            return Array(body, pos)
        if k == lx.LCURLY:
            self._expect(lx.LCURLY)
            body = self._list()
            self._expect(lx.RCURLY)
            return Function(body, pos)
        self._take()
        if k == lx.IDENT:
            return Identifier(t.text, self.ids.get_or_create(t.text), pos)
        if k == lx.BINDER:
            name = t.text[1:]
            return Binder(name, self.ids.get_or_create(name), pos)
        if k == lx.INT:
            try:
                v = int(t.text, 10)
                if not (-(1 << 63) <= v < (1 << 63)):
                    raise ValueError
            except ValueError:
                raise GMLSyntaxError(f"{t.line}:{t.col}: could not parse number: {t.text}")
            return IntLiteral(v, pos)
        if k == lx.FLOAT:
            try:
                v = float(t.text)
            except ValueError:
                raise GMLSyntaxError(f"{t.line}:{t.col}: could not parse number: {t.text}")
            return FloatLiteral(v, pos)
        if k == lx.BOOLEAN:
            return BoolLiteral(t.text == "true", pos)
        if k == lx.STRING:
            return StringLiteral(t.text, pos)
        raise GMLSyntaxError(f"{t.line}:{t.col}: unexpected token: {k}")
