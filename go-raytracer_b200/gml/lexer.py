"""GML tokeniser + mini preprocessor.

Host-side mirror of the reference's text front end
(/root/reference/internal/gml/lexer.go:170-230 token rules, :264-386 directives,
:388-477 identifier / number / string scanning).  It exists because the Go tool
chain is absent in this build environment; it is NOT on the accelerated path.

Implementation note: unlike the reference (a byte-at-a-time state machine with a
frame stack), this is a generator over an explicit stack of ``_Source`` cursors;
the observable token stream and the error strings are the same.
"""
import os

# token kinds (names follow lexer.go:41-56 so parser errors read the same)
EOF, ILLEGAL, ERROR, IDENT, BINDER, BOOLEAN, INT, FLOAT, STRING, LCURLY, RCURLY, LBRACKET, RBRACKET = (
    "EOF", "Illegal", "Error", "Ident", "Binder", "Boolean", "Integer", "Float", "String",
    "LCurly", "RCurly", "LBracket", "RBracket")

_PUNCT = {"{": LCURLY, "}": RCURLY, "[": LBRACKET, "]": RBRACKET}
_LETTERS = frozenset("abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ")
_DIGITS = frozenset("0123456789")
_IDENT_TAIL = _LETTERS | _DIGITS | frozenset("-_")


class Token:
    __slots__ = ("kind", "text", "line", "col")

    def __init__(self, kind, text, line, col):
        self.kind, self.text, self.line, self.col = kind, text, line, col

    def __repr__(self):
        return f"Token({self.kind}, {self.text!r}, {self.line}:{self.col})"


class _Source:
    """One input buffer (the top-level text or one #include'd file)."""
    __slots__ = ("text", "i", "line", "col", "path")

    def __init__(self, text, path=""):
        self.text, self.i, self.line, self.col, self.path = text, 0, 1, 1, path


class Lexer:
    def __init__(self, text, path=""):
        path = os.path.abspath(path) if path else ""
        self._stack = [_Source(text, path)]
        self._open = {path} if path else set()   # files being read, for cycle detection
        self._defined = set()                    # names seen in #define
        self._cond_depth = 0

    @classmethod
    def from_file(cls, path):
        with open(path, "r", encoding="utf-8", errors="surrogateescape") as f:
            return cls(f.read(), path)

    # --- character cursor -------------------------------------------------
    def _peek(self, k=0):
        s = self._stack[-1]
        j = s.i + k
        return s.text[j] if j < len(s.text) else ""

    def _advance(self):
        s = self._stack[-1]
        if s.i < len(s.text):
            if s.text[s.i] == "\n":
                s.line += 1
                s.col = 1
            else:
                s.col += 1
            s.i += 1

    def _at_end_pop(self):
        """At end of an included file resume the includer (lexer.go:131-151)."""
        while len(self._stack) > 1 and self._stack[-1].i >= len(self._stack[-1].text):
            done = self._stack.pop()
            self._open.discard(done.path)

    def _pos(self):
        s = self._stack[-1]
        return s.line, s.col

    # --- public -----------------------------------------------------------
    def next_token(self):
        while True:
            self._at_end_pop()
            c = self._peek()
            if c in (" ", "\t", "\n", "\r") and c:
                self._advance()
                continue
            line, col = self._pos()
            if c == "":
                return Token(EOF, "", line, col)
            if c in _PUNCT:
                self._advance()
                return Token(_PUNCT[c], c, line, col)
            if c == "%":
                while self._peek() not in ("\n", ""):
                    self._advance()
                continue
            if c == "/":
                n = self._peek(1)
                if n in _LETTERS and n:
                    self._advance()
                    return Token(BINDER, "/" + self._ident(), line, col)
                if n == "*":
                    err = self._block_comment()
                    if err:
                        return Token(ERROR, err, line, col)
                    continue
                self._advance()
                return Token(ILLEGAL, "/", line, col)
            if c == '"':
                text, err = self._string()
                return Token(ILLEGAL if err else STRING, text, line, col)
            if c == "#":
                err = self._directive()
                if err:
                    return Token(ERROR, err, line, col)
                continue
            if c in _LETTERS:
                word = self._ident()
                return Token(BOOLEAN if word in ("true", "false") else IDENT, word, line, col)
            if c in _DIGITS or c == "-":
                text, kind = self._number()
                return Token(kind, text, line, col)
            self._advance()
            return Token(ILLEGAL, c, line, col)

    # --- scanners -----------------------------------------------------------
    def _ident(self):
        s = self._stack[-1]
        j = s.i
        while self._peek() in _IDENT_TAIL and self._peek():
            self._advance()
        return s.text[j:s.i]

    def _number(self):
        s = self._stack[-1]
        j = s.i
        kind = INT
        if self._peek() == "-":
            self._advance()
        while self._peek() in _DIGITS and self._peek():
            self._advance()
        if self._peek() == ".":
            kind = FLOAT
            self._advance()
            while self._peek() in _DIGITS and self._peek():
                self._advance()
        if self._peek() in ("e", "E") and self._peek():
            kind = FLOAT
            self._advance()
            if self._peek() in ("+", "-") and self._peek():
                self._advance()
            while self._peek() in _DIGITS and self._peek():
                self._advance()
        return s.text[j:s.i], kind

    def _string(self):
        out, err = [], None
        self._advance()  # opening quote
        while self._peek() not in ('"', ""):
            c = self._peek()
            if c == "\\":
                self._advance()
                e = self._peek()
                if e in ("n", "t", '"', "\\") and e:
                    out.append({"n": "\n", "t": "\t", '"': '"', "\\": "\\"}[e])
                else:
                    err = "illegal escape sequence"
                    out.append("\\" + e)
            else:
                out.append(c)
            self._advance()
        if self._peek() == '"':
            self._advance()
        elif err is None:
            err = "unclosed string literal"
        return "".join(out), err

    def _block_comment(self):
        self._advance()
        self._advance()
        while True:
            c = self._peek()
            if c == "":
                return "unterminated block comment"
            if c == "*" and self._peek(1) == "/":
                self._advance()
                self._advance()
                return None
            self._advance()

    def _inline_space(self):
        while self._peek() in (" ", "\t") and self._peek():
            self._advance()

    def _directive(self):
        self._advance()  # '#'
        self._inline_space()
        word = self._ident()
        if word == "include":
            self._inline_space()
            if self._peek() != '"':
                return "expected quoted filename after #include"
            name, err = self._string()
            if err:
                return f"invalid #include filename: {err}"
            return self._push_include(name)
        if word == "ifndef":
            self._inline_space()
            name = self._ident()
            if not name:
                return "expected identifier after #ifndef"
            if name in self._defined:
                return self._skip_conditional()
            self._cond_depth += 1
            return None
        if word == "define":
            self._inline_space()
            name = self._ident()
            if not name:
                return "expected identifier after #define"
            self._defined.add(name)
            return None
        if word == "endif":
            if self._cond_depth == 0:
                return "#endif without matching #ifndef"
            self._cond_depth -= 1
            return None
        return f"unsupported preprocessor directive: #{word}"

    def _push_include(self, name):
        here = self._stack[-1].path
        base = os.path.dirname(here) if here else "."
        path = os.path.abspath(os.path.join(base, name))
        try:
            with open(path, "r", encoding="utf-8", errors="surrogateescape") as f:
                text = f.read()
        except OSError as e:
            return f'#include "{name}": {e.strerror or e}'
        if path in self._open:
            return f'#include "{name}": include cycle detected'
        self._open.add(path)
        self._stack.append(_Source(text, path))
        return None

    def _skip_conditional(self):
        depth = 1
        while depth > 0:
            c = self._peek()
            if c == "":
                return "unterminated #ifndef: missing #endif"
            if c == "#":
                self._advance()
                self._inline_space()
                w = self._ident()
                if w == "ifndef":
                    depth += 1
                elif w == "endif":
                    depth -= 1
                continue
            self._advance()
        return None
