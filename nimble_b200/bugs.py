"""A small parser for the subset of the BUGS language the supported hot path needs.

The reference parses model code with R's own parser (`nimbleCode` is `quote`) and turns every line into a
BUGSdeclClass (P/R/BUGS_BUGSdecl.R:111-170) inside nested for-loop contexts (P/R/BUGS_contexts.R).  R is not
available to the host layer here, so this module produces the same information from text:
statements `lhs ~ dist(args)`, `lhs <- expr`, `link(lhs) <- expr`, nested `for (i in a:b) { ... }`.

AST nodes are tuples:
  ('num', v)  ('var', name, [index, ...])  ('range', lo, hi)  ('call', fname, [args], {kw: arg})
  ('bin', op, l, r)  ('neg', e)  ('sub', expr, [index, ...])   (op may be '%*%')
Statements: ('for', var, lo, hi, [stmts])  ('stoch', lhs, dist, args, kwargs)  ('determ', link|None, lhs, expr)
"""
import re

_TOK = re.compile(r"\s*(?:(\d+\.\d*(?:[eE][-+]?\d+)?|\.\d+(?:[eE][-+]?\d+)?|\d+(?:[eE][-+]?\d+)?)|([A-Za-z_.][A-Za-z0-9_.]*)|(<-|%\*%|[-+*/^()\[\]{},:~=;]))")


def _tokenize(text):
    text = re.sub(r"#[^\n]*", "", text)
    toks, pos = [], 0
    text = text.replace("\n", " ; ")
    while pos < len(text):
        if text[pos:].strip() == "":
            break
        m = _TOK.match(text, pos)
        if not m:
            raise SyntaxError("BUGS code: cannot tokenize near %r" % text[pos:pos + 30])
        if m.group(1) is not None:
            toks.append(("num", float(m.group(1))))
        elif m.group(2) is not None:
            toks.append(("id", m.group(2)))
        else:
            toks.append(("op", m.group(3)))
        pos = m.end()
    return toks


class _P:
    def __init__(self, toks):
        self.t, self.i = toks, 0

    def peek(self, k=0):
        return self.t[self.i + k] if self.i + k < len(self.t) else ("eof", None)

    def next(self):
        tok = self.peek()
        self.i += 1
        return tok

    def accept(self, v):
        if self.peek() == ("op", v):
            self.i += 1
            return True
        return False

    def expect(self, v):
        if not self.accept(v):
            raise SyntaxError("BUGS code: expected %r, found %r" % (v, self.peek()))

    def skip_sep(self):
        while self.accept(";"):
            pass

    # expressions: + -  <  * /  <  unary -  <  ^
    def expr(self):
        e = self.term()
        while self.peek() in (("op", "+"), ("op", "-")):
            op = self.next()[1]
            e = ("bin", op, e, self.term())
        return e

    def term(self):
        e = self.unary()
        while self.peek() in (("op", "*"), ("op", "/"), ("op", "%*%")):
            op = self.next()[1]
            e = ("bin", op, e, self.unary())
        return e

    def unary(self):
        if self.accept("-"):
            return ("neg", self.unary())
        if self.accept("+"):
            return self.unary()
        return self.power()

    def power(self):
        b = self.atom()
        if self.accept("^"):
            return ("bin", "^", b, self.unary())
        return b

    def index(self):
        e = self.expr()
        if self.accept(":"):
            return ("range", e, self.expr())
        return e

    def atom(self):
        k, v = self.next()
        if k == "num":
            return ("num", v)
        if k == "op" and v == "(":
            e = self.expr()
            self.expect(")")
            if self.accept("["):   # (x[i, 1:p] %*% beta[1:p])[1, 1]: an index applied to an expression
                idx = []
                while not self.accept("]"):
                    idx.append(self.index())
                    self.accept(",")
                return ("sub", e, idx)
            return e
        if k == "id":
            if self.accept("("):
                args, kw = [], {}
                while not self.accept(")"):
                    if self.peek()[0] == "id" and self.peek(1) == ("op", "="):
                        name = self.next()[1]
                        self.next()
                        kw[name] = self.index()
                    else:
                        args.append(self.index())
                    self.accept(",")
                return ("call", v, args, kw)
            idx = []
            if self.accept("["):
                while not self.accept("]"):
                    idx.append(self.index())
                    self.accept(",")
            return ("var", v, idx)
        raise SyntaxError("BUGS code: unexpected token %r" % ((k, v),))

    def block(self):
        stmts = []
        self.skip_sep()
        while self.peek()[0] != "eof" and self.peek() != ("op", "}"):
            stmts.append(self.stmt())
            self.skip_sep()
        return stmts

    def stmt(self):
        if self.peek() == ("id", "for"):
            self.next()
            self.expect("(")
            var = self.next()[1]
            if self.next() != ("id", "in"):
                raise SyntaxError("BUGS code: expected 'in'")
            lo = self.expr()
            self.expect(":")
            hi = self.expr()
            self.expect(")")
            self.skip_sep()
            self.expect("{")
            body = self.block()
            self.expect("}")
            return ("for", var, lo, hi, body)
        lhs = self.atom()
        if self.accept("~"):
            call = self.atom()
            if call[0] != "call":
                raise SyntaxError("BUGS code: expected a distribution call after '~'")
            return ("stoch", lhs, call[1], call[2], call[3])
        self.expect("<-")
        rhs = self.expr()
        if lhs[0] == "call":  # link(lhs) <- expr
            return ("determ", lhs[1], lhs[2][0], rhs)
        return ("determ", None, lhs, rhs)


def parse(text):
    """Parse BUGS model text (with or without the surrounding `model { }`) into a statement list."""
    text = text.strip()
    m = re.match(r"^(?:model|nimbleCode\s*\()?\s*\{(.*)\}\s*\)?\s*$", text, re.S)
    if m:
        text = m.group(1)
    p = _P(_tokenize(text))
    stmts = p.block()
    if p.peek()[0] != "eof":
        raise SyntaxError("BUGS code: trailing tokens at %r" % (p.peek(),))
    return stmts
