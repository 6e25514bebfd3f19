"""TEST INFRASTRUCTURE ONLY -- never imported by the product path.

A small interpreter for the subset of MATLAB that the reference's filter classes are written in (classdef with
property defaults and methods, functions with several outputs, if / for / switch, matrix literals with MATLAB's
white-space rules, ranges, 1-based indexing with growth on assignment, transpose, the matrix operators, struct-array
field lists `[a.f]`, handle and value classes).  It exists for one purpose: to EXECUTE THE UNMODIFIED SOURCE TEXT of

    EKF_SLAM.m   EKF_SLAM_UC.m   Correspondence.m   (classes)      append.m   example.m: predict, f   (free functions)

from a checkout of SamShue/EKF_SLAM (read where it lies, never copied) on the committed golden inputs, so that
oracle/dense_literal.py -- the hand restatement every parity test leans on -- is compared with what the reference's own
statements compute, statement by statement as MATLAB's grammar orders them, and not only with a reading of them
(tests/golden/make_reference_run.py writes the vectors, tests/test_oracle_cpu.py checks the oracle against them).

A second use (tests/matlab_host.py, tests/test_matlab_classes.py): the repo's own MATLAB host classes in matlab/ are
executed by the same interpreter -- for them it also knows superclass chains (`h@Base(...)`, `method@Base(h, ...)`),
Dependent properties served by get.p / set.p, nargin, scalar structs, `~` inputs and stand-ins for MEX files that
see nargout -- because no MATLAB or Octave exists here to run them.

What this does and does not pin.  MATLAB and Octave are absent from this image and from the GPU box; the MathWorks
run-time functions the reference calls (sind, cosd, atan2d, wrapTo360, the LAPACK inverse behind `^-1`, the BLAS behind
`*`) are closed source and are served here by the same restatements the oracle uses (oracle/matlab_compat.py, NumPy).  So a
run through this interpreter pins the TRANSLATION of the reference -- control flow, operator precedence and
associativity, indexing, array growth, the quirks (idx = ii, the un-wrapped innovation, the signature-only score) --
to the reference text itself, mechanically; it does not pin MathWorks' built-ins.  "Parity unpinned" therefore stays
in DESIGN.md for the latter, and matlab/run_reference_fixture.m remains the witness for whoever has MATLAB or Octave.

Grammar notes (only what the three files need; anything else raises MiniMatlabError rather than guessing):
  * `'` is a transpose directly after an operand, a string delimiter otherwise;
  * inside `[ ]`, white space separates elements, and `a -b` is two elements while `a - b` and `a-b` are one;
  * precedence, tightest first: postfix ' and ^ ; unary + - ~ ; * / .* ./ ; + - ; : ; comparisons ; & ; | ; && ; ||
    (an exponent may carry its own unary minus: `phi_k^-1`);
  * `x(i)` on a vector keeps the vector's orientation; `x(i) = v`, `A(i,j) = V` grow the array with zeros;
  * `A^-1` of a square matrix is the inverse (matlab_compat.inv2), `a*B` with a scalar operand is element-wise;
  * `name word` at statement level is command syntax, name('word')  (`hold on` in the classes' plot method);
  * a file that is not a classdef contributes its functions (append.m; the local functions behind example.m's script part).
"""
import copy
import re

import numpy as np

from . import matlab_compat as MC


class MiniMatlabError(RuntimeError):
    pass


# ------------------------------------------------------------------------------------------------------ tokens
_KEYWORDS = {'if', 'elseif', 'else', 'end', 'for', 'while', 'switch', 'case', 'otherwise', 'function', 'classdef',
             'properties', 'methods', 'return', 'break', 'continue'}
_NUM = re.compile(r'(\d+\.?\d*([eE][+-]?\d+)?|\.\d+([eE][+-]?\d+)?)')
_ID = re.compile(r'[A-Za-z_]\w*')
_OPS = ['...', '==', '~=', '<=', '>=', '&&', '||', '.*', './', '.^', ".'", '+', '-', '*', '/', '^', '<', '>', '=', '~',
        '&', '|', ':', '(', ')', '[', ']', ',', ';', '.', '@', '{', '}']


class Tok:
    __slots__ = ('kind', 'val', 'sp', 'line')

    def __init__(self, kind, val, sp, line):
        self.kind, self.val, self.sp, self.line = kind, val, sp, line

    def __repr__(self):
        return '%s(%r)@%d' % (self.kind, self.val, self.line)


def tokenize(src):
    toks = []
    for ln, line in enumerate(src.splitlines(), 1):
        i, n = 0, len(line)
        sp = True
        cont = False
        while i < n:
            c = line[i]
            if c in ' \t':
                i += 1; sp = True
                continue
            if c == '%':
                break
            if line.startswith('...', i):
                cont = True
                break
            if c == "'":
                prev = toks[-1] if toks else None
                is_transpose = (prev is not None and not sp and prev.kind != 'NL' and
                                (prev.kind in ('NUM', 'ID') or prev.val in (')', ']', "'", ".'") or
                                 (prev.kind == 'KW' and prev.val == 'end')))
                if is_transpose:
                    toks.append(Tok('OP', "'", sp, ln)); i += 1; sp = False
                    continue
                j, out = i + 1, []
                while True:
                    if j >= n:
                        raise MiniMatlabError('line %d: unterminated string' % ln)
                    if line[j] == "'":
                        if j + 1 < n and line[j + 1] == "'":
                            out.append("'"); j += 2
                            continue
                        break
                    out.append(line[j]); j += 1
                toks.append(Tok('STR', ''.join(out), sp, ln)); i = j + 1; sp = False
                continue
            m = _NUM.match(line, i)
            if m and (c.isdigit() or (c == '.' and i + 1 < n and line[i + 1].isdigit())):
                toks.append(Tok('NUM', float(m.group(0)), sp, ln)); i = m.end(); sp = False
                continue
            m = _ID.match(line, i)
            if m:
                w = m.group(0)
                toks.append(Tok('KW' if w in _KEYWORDS else 'ID', w, sp, ln)); i = m.end(); sp = False
                continue
            for op in _OPS:
                if line.startswith(op, i):
                    toks.append(Tok('OP', op, sp, ln)); i += len(op); sp = False
                    break
            else:
                raise MiniMatlabError('line %d: cannot tokenize %r' % (ln, line[i:i + 10]))
        if not cont:
            toks.append(Tok('NL', '\n', True, ln))
    toks.append(Tok('EOF', None, True, -1))
    return toks


# ------------------------------------------------------------------------------------------------------ parser
class Parser:
    def __init__(self, toks):
        self.t, self.i = toks, 0
        self.mdepth = 0       # > 0: directly inside a matrix literal (white space separates elements)

    def peek(self, k=0):
        return self.t[self.i + k]

    def next(self):
        tok = self.t[self.i]
        self.i += 1
        return tok

    def at(self, val, kind=None):
        tok = self.peek()
        return tok.val == val and (kind is None or tok.kind == kind) and tok.kind not in ('STR',)

    def expect(self, val):
        tok = self.next()
        if tok.val != val or tok.kind == 'STR':
            raise MiniMatlabError('line %d: expected %r, got %r' % (tok.line, val, tok.val))
        return tok

    def skip_nl(self):
        while self.peek().kind == 'NL' or (self.peek().kind == 'OP' and self.peek().val in (';', ',')):
            self.next()

    # ---- file level
    def parse_file(self):
        self.skip_nl()
        if self.peek().kind == 'KW' and self.peek().val == 'classdef':
            return self.parse_classdef()
        return self.parse_functions()

    def parse_functions(self):
        """A function file (append.m), or a script whose local functions follow the script part (example.m): the script
        statements are skipped token by token up to the first `function` that starts a line; only the functions are kept."""
        funcs = {}
        at_line_start = True
        while self.peek().kind != 'EOF':
            tok = self.peek()
            if at_line_start and tok.kind == 'KW' and tok.val == 'function':
                f = self.parse_function()
                funcs[f['name']] = f
                at_line_start = True
                continue
            self.next()
            at_line_start = tok.kind == 'NL'
        if not funcs:
            raise MiniMatlabError('no classdef and no function in this file')
        return {'kind': 'functions', 'functions': funcs}

    def parse_classdef(self):
        self.expect('classdef')
        name = self.next().val
        sup = None
        if self.at('<'):
            self.next(); sup = self.next().val
        props, methods, dependent = [], {}, set()
        self.skip_nl()
        while not (self.peek().kind == 'KW' and self.peek().val == 'end'):
            kw = self.next()
            if kw.val == 'properties':
                attrs = self._skip_attrs(); self.skip_nl()
                while not (self.peek().kind == 'KW' and self.peek().val == 'end'):
                    pname = self.next()
                    if pname.kind != 'ID':
                        raise MiniMatlabError('line %d: property name expected' % pname.line)
                    default = None
                    if self.at('='):
                        self.next(); default = self.parse_expr()
                    if 'Dependent' in attrs:
                        dependent.add(pname.val)         # no storage: served by get.<name> / set.<name>
                    else:
                        props.append((pname.val, default))
                    self.skip_nl()
                self.expect('end'); self.skip_nl()
            elif kw.val == 'methods':
                self._skip_attrs(); self.skip_nl()
                while not (self.peek().kind == 'KW' and self.peek().val == 'end'):
                    f = self.parse_function()
                    methods[f['name']] = f
                    self.skip_nl()
                self.expect('end'); self.skip_nl()
            else:
                raise MiniMatlabError('line %d: unexpected %r in classdef' % (kw.line, kw.val))
        self.expect('end')
        return {'kind': 'class', 'name': name, 'super': sup, 'props': props, 'methods': methods, 'dependent': dependent}

    def _skip_attrs(self):
        seen = []
        if self.at('('):
            depth = 0
            while True:
                tok = self.next()
                if tok.val == '(' and tok.kind == 'OP':
                    depth += 1
                elif tok.val == ')' and tok.kind == 'OP':
                    depth -= 1
                    if depth == 0:
                        break
                elif tok.kind == 'ID':
                    seen.append(tok.val)
        return seen

    def parse_function(self):
        self.expect('function')
        outs = []
        if self.at('['):
            self.next()
            while not self.at(']'):
                tok = self.next()
                if tok.kind == 'ID':
                    outs.append(tok.val)
            self.expect(']'); self.expect('=')
            name = self.next().val
        else:
            first = self.next().val
            if self.at('='):
                self.next(); outs = [first]; name = self.next().val
            else:
                name = first
        if name in ('get', 'set') and self.at('.') and self.peek(1).kind == 'ID':      # property accessors: get.x / set.x
            self.next(); name = name + '.' + self.next().val
        params = []
        if self.at('('):
            self.next()
            while not self.at(')'):
                tok = self.next()
                if tok.kind == 'ID':
                    params.append(tok.val)
                elif tok.kind == 'OP' and tok.val == '~':
                    params.append('~')                     # an ignored input still takes its position
            self.expect(')')
        body = self.parse_block(('end',))
        self.expect('end')
        return {'kind': 'function', 'name': name, 'outs': outs, 'params': params, 'body': body}

    # ---- statements
    def parse_block(self, terminators):
        out = []
        self.skip_nl()
        while not (self.peek().kind == 'KW' and self.peek().val in terminators):
            if self.peek().kind == 'EOF':
                raise MiniMatlabError('unexpected end of file')
            out.append(self.parse_statement())
            self.skip_nl()
        return out

    def parse_statement(self):
        tok = self.peek()
        if tok.kind == 'KW':
            if tok.val == 'if':
                self.next()
                clauses, other = [], None
                cond = self.parse_expr()
                body = self.parse_block(('elseif', 'else', 'end'))
                clauses.append((cond, body))
                while True:
                    kw = self.next()
                    if kw.val == 'elseif':
                        cond = self.parse_expr(); clauses.append((cond, self.parse_block(('elseif', 'else', 'end'))))
                    elif kw.val == 'else':
                        other = self.parse_block(('end',))
                    else:
                        break
                return ('if', clauses, other, tok.line)
            if tok.val == 'for':
                self.next()
                paren = self.at('(')
                if paren:
                    self.next()
                var = self.next().val
                self.expect('=')
                rng = self.parse_expr()
                if paren:
                    self.expect(')')
                body = self.parse_block(('end',))
                self.expect('end')
                return ('for', var, rng, body, tok.line)
            if tok.val == 'switch':
                self.next()
                subj = self.parse_expr()
                self.skip_nl()
                cases, other = [], None
                while not (self.peek().kind == 'KW' and self.peek().val == 'end'):
                    kw = self.next()
                    if kw.val == 'case':
                        val = self.parse_expr(); cases.append((val, self.parse_block(('case', 'otherwise', 'end'))))
                    elif kw.val == 'otherwise':
                        other = self.parse_block(('case', 'otherwise', 'end'))
                    else:
                        raise MiniMatlabError('line %d: unexpected %r in switch' % (kw.line, kw.val))
                self.expect('end')
                return ('switch', subj, cases, other, tok.line)
            if tok.val == 'return':
                self.next()
                return ('return', tok.line)
            if tok.val == 'while':
                self.next()
                cond = self.parse_expr()
                body = self.parse_block(('end',))
                self.expect('end')
                return ('while', cond, body, tok.line)
            if tok.val in ('break', 'continue'):
                self.next()
                return (tok.val, tok.line)
            raise MiniMatlabError('line %d: statement %r not supported' % (tok.line, tok.val))
        if tok.kind == 'ID' and self.peek(1).kind == 'ID' and self.peek(1).sp and \
                (self.peek(2).kind in ('NL', 'EOF') or (self.peek(2).kind == 'OP' and self.peek(2).val in (';', ','))):
            self.next(); arg = self.next().val              # command syntax: `hold on` is hold('on')
            return ('expr', ('index', ('id', tok.val), [('str', arg)]), tok.line)
        lhs = self.parse_expr()
        if self.at('='):
            self.next()
            rhs = self.parse_expr()
            if lhs[0] == 'matrix':
                if len(lhs[1]) != 1:
                    raise MiniMatlabError('line %d: bad multiple assignment' % tok.line)
                return ('massign', lhs[1][0], rhs, tok.line)
            return ('assign', lhs, rhs, tok.line)
        return ('expr', lhs, tok.line)

    # ---- expressions (precedence climbing, loosest first)
    def parse_expr(self):
        return self._oror()

    def _binop_loop(self, sub, ops):
        a = sub()
        while self.peek().kind == 'OP' and self.peek().val in ops and not self._elem_break():
            op = self.next().val
            a = ('bin', op, a, sub())
        return a

    def _elem_break(self):
        """Inside a matrix literal `a -b` / `a +b` starts a new element (space before the sign, none after it)."""
        if self.mdepth == 0:
            return False
        tok, nxt = self.peek(), self.peek(1)
        return tok.val in ('+', '-') and tok.sp and not nxt.sp

    def _oror(self):
        return self._binop_loop(self._andand, ('||',))

    def _andand(self):
        return self._binop_loop(self._or, ('&&',))

    def _or(self):
        return self._binop_loop(self._and, ('|',))

    def _and(self):
        return self._binop_loop(self._cmp, ('&',))

    def _cmp(self):
        return self._binop_loop(self._range, ('==', '~=', '<', '<=', '>', '>='))

    def _range(self):
        a = self._add()
        if self.at(':') and self.peek().kind == 'OP' and not self._colon_is_index_all():
            self.next()
            b = self._add()
            if self.at(':') and self.peek().kind == 'OP':
                self.next()
                c = self._add()
                return ('range', a, c, b)          # a:step:c
            return ('range', a, b, None)
        return a

    def _colon_is_index_all(self):
        return False

    def _add(self):
        return self._binop_loop(self._mul, ('+', '-'))

    def _mul(self):
        return self._binop_loop(self._unary, ('*', '/', '.*', './'))

    def _unary(self):
        tok = self.peek()
        if tok.kind == 'OP' and tok.val in ('+', '-', '~'):
            self.next()
            return ('un', tok.val, self._unary())
        return self._power()

    def _power(self):
        a = self._postfix()
        while self.peek().kind == 'OP' and self.peek().val in ('^', '.^'):
            op = self.next().val
            nt = self.peek()
            if nt.kind == 'OP' and nt.val in ('+', '-', '~'):       # 2^-1: the exponent carries its own sign
                self.next()
                b = ('un', nt.val, self._postfix())
            else:
                b = self._postfix()
            a = ('bin', op, a, b)
        return a

    def _postfix(self):
        a = self._primary()
        while True:
            tok = self.peek()
            if tok.kind == 'OP' and tok.val in ("'", ".'"):
                self.next(); a = ('transpose', a)
            elif tok.kind == 'OP' and tok.val == '(' and not (self.mdepth > 0 and tok.sp):
                self.next()
                saved, self.mdepth = self.mdepth, 0
                self.idepth = getattr(self, 'idepth', 0) + 1
                args = []
                while not self.at(')'):
                    if self.at(':') and self.peek(1).val in (',', ')'):
                        self.next(); args.append(('colon_all',))
                    else:
                        args.append(self.parse_expr())
                    if self.at(','):
                        self.next()
                self.expect(')')
                self.mdepth = saved
                self.idepth -= 1
                a = ('index', a, args)
            elif tok.kind == 'OP' and tok.val == '.' and self.peek(1).kind == 'ID' and not tok.sp:
                self.next(); a = ('field', a, self.next().val)
            elif tok.kind == 'OP' and tok.val == '.' and self.peek(1).kind == 'OP' and self.peek(1).val == '(' and not tok.sp:
                self.next(); self.next()                    # o.(name): the field whose name is the value of an expression
                saved, self.mdepth = self.mdepth, 0
                e = self.parse_expr()
                self.expect(')')
                self.mdepth = saved
                a = ('dynfield', a, e)
            elif tok.kind == 'OP' and tok.val == '@' and not tok.sp and a[0] == 'id' and self.peek(1).kind == 'ID':
                self.next(); a = ('superref', a[1], self.next().val)     # h@Base(...) / method@Base(h, ...)
            else:
                return a

    def _primary(self):
        tok = self.next()
        if tok.kind == 'NUM':
            return ('num', tok.val)
        if tok.kind == 'STR':
            return ('str', tok.val)
        if tok.kind == 'ID':
            return ('id', tok.val)
        if tok.kind == 'KW' and tok.val == 'end' and getattr(self, 'idepth', 0) > 0:
            return ('end',)                                 # last index of the dimension it stands in
        if tok.kind == 'OP' and tok.val == '(':
            saved, self.mdepth = self.mdepth, 0
            e = self.parse_expr()
            self.expect(')')
            self.mdepth = saved
            return ('paren', e)
        if tok.kind == 'OP' and tok.val == '{':
            self.expect('}')                                # only the empty cell, as in struct('f', {}) -> 0x0 struct array
            return ('emptycell',)
        if tok.kind == 'OP' and tok.val == '[':
            saved, self.mdepth = self.mdepth, 1
            rows, row = [], []
            while True:
                nt = self.peek()
                if nt.kind == 'OP' and nt.val == ']':
                    self.next()
                    break
                if nt.kind == 'NL' or (nt.kind == 'OP' and nt.val == ';'):
                    self.next()
                    if row:
                        rows.append(row); row = []
                    continue
                if nt.kind == 'OP' and nt.val == ',':
                    self.next()
                    continue
                if nt.kind == 'OP' and nt.val == '~' and self.peek(1).kind == 'OP' and self.peek(1).val in (',', ']'):
                    self.next(); row.append(('tilde',))     # [~, order] = sort(v): an output nobody keeps
                    continue
                row.append(self.parse_expr())
            if row:
                rows.append(row)
            self.mdepth = saved
            return ('matrix', rows)
        raise MiniMatlabError('line %d: unexpected token %r' % (tok.line, tok.val))


# ------------------------------------------------------------------------------------------------------ values
class MObject:
    def __init__(self, cls):
        self.cls = cls
        self.props = {}

    def copy(self):
        o = MObject(self.cls)
        o.props = {k: _copy_value(v) for k, v in self.props.items()}
        return o


class BoundMethod:
    def __init__(self, obj, func):
        self.obj, self.func = obj, func


class StructArray:
    """A MATLAB struct array backed by Python objects with attributes (the landmark table handed in by the tests)."""

    def __init__(self, items):
        self.items = list(items)


class EmptyCell:
    pass


class CSList:
    def __init__(self, vals):
        self.vals = vals


class PyObj:
    """A Python object standing in for a MATLAB object the reference receives from outside (landmark_list)."""

    def __init__(self, o):
        self.o = o


class MStruct:
    """A scalar MATLAB struct built inside the interpreter (struct(), s.f = v): a value type."""

    def __init__(self, fields=None):
        self.f = dict(fields or {})


class _Return(Exception):
    pass


class _Break(Exception):
    pass


class _Continue(Exception):
    pass


def _copy_value(v):
    if isinstance(v, np.ndarray):
        return v.copy()
    if isinstance(v, MObject):
        return v if v.cls['handle'] else v.copy()
    if isinstance(v, MStruct):
        return MStruct({k: _copy_value(e) for k, e in v.f.items()})
    return v


def to_matlab(v):
    """Python value -> interpreter value."""
    if v is None:
        return np.zeros((0, 0))
    if isinstance(v, (MObject, StructArray, PyObj, str, BoundMethod, MStruct, EmptyCell)):
        return v
    if isinstance(v, np.ndarray) and v.dtype == np.uint64:
        return v.reshape(1, 1).copy() if v.ndim < 2 else v.copy()
    if isinstance(v, (bool, np.bool_)):
        return np.array([[bool(v)]])
    if isinstance(v, (int, float, np.integer, np.floating)):
        return np.array([[float(v)]])
    if isinstance(v, np.ndarray):
        a = v.astype(np.float64) if v.dtype != np.bool_ else v
        if a.ndim == 0:
            return a.reshape(1, 1)
        if a.ndim == 1:
            return a.reshape(1, -1).copy()
        return a.copy()
    if isinstance(v, (list, tuple)):
        if len(v) and all(isinstance(e, (int, float, np.integer, np.floating)) for e in v):
            return np.array([list(map(float, v))])
        if len(v) == 0 or hasattr(v[0], '__dict__'):
            return StructArray(v)
        return to_matlab(np.asarray(v, dtype=np.float64))
    if hasattr(v, '__dict__'):
        return PyObj(v)
    raise MiniMatlabError('cannot bring %r into the interpreter' % (type(v),))


def _is_scalar(a):
    return isinstance(a, np.ndarray) and a.size == 1


def _truth(v):
    if isinstance(v, np.ndarray):
        return v.size > 0 and bool(np.all(v != 0))
    raise MiniMatlabError('condition is not numeric')


def _num(v):
    if not isinstance(v, np.ndarray):
        raise MiniMatlabError('numeric value expected, got %r' % (type(v),))
    return v.astype(np.float64) if v.dtype == np.bool_ else v


# ------------------------------------------------------------------------------------------------------ interpreter
class Interpreter:
    def __init__(self, stubs=None):
        self.classes = {}
        self._end_stack = []
        self.path, self._loaded, self._files, self._fids = [], set(), [], {}
        self._in_accessor = set()        # (object, accessor) pairs being served: inside get.p / set.p the property is plain
        self.functions = {}              # function files / local functions of a script, by name
        self.builtins = {
            'length': lambda a: float(max(a.shape)) if isinstance(a, np.ndarray) and a.size else 0.0,
            'numel': lambda a: float(len(a.items)) if isinstance(a, StructArray) else float(a.size),
            'isempty': lambda a: (a.size == 0) if isinstance(a, np.ndarray) else
                                 (len(a.items) == 0 if isinstance(a, StructArray) else (len(a) == 0 if isinstance(a, str) else False)),
            'sqrt': lambda a: np.sqrt(_num(a)),
            'abs': lambda a: np.abs(_num(a)),
            'cosd': lambda a: self._ew(MC.cosd, a),
            'sind': lambda a: self._ew(MC.sind, a),
            'atan2d': lambda y, x: np.array([[MC.atan2d(float(_num(y).ravel()[0]), float(_num(x).ravel()[0]))]])
                                   if _num(y).size == 1 and _num(x).size == 1 else
                                   np.vectorize(MC.atan2d, otypes=[np.float64])(_num(y), _num(x)),
            'wrapTo360': lambda a: self._ew(MC.wrapTo360, a),
            'cos': lambda a: np.cos(_num(a)),
            'sin': lambda a: np.sin(_num(a)),
            'atan': lambda a: np.arctan(_num(a)),
            'eig': None,                 # two outputs: handled in call()
            'uint64': lambda a: np.array([[int(_num(a).ravel()[0])]], dtype=np.uint64),
            'strcmp': lambda a, b: isinstance(a, str) and isinstance(b, str) and a == b,
            'max': lambda a, b=None: np.max(_num(a)) if b is None else np.maximum(_num(a), _num(b)),
            'min': lambda a, b=None: np.min(_num(a)) if b is None else np.minimum(_num(a), _num(b)),
            'struct': self._struct,
            'mod': lambda a, b: np.mod(_num(a), _num(b)),
            'sum': lambda a, d=None: (np.sum(_num(a)) if min(_num(a).shape) == 1 else np.sum(_num(a), axis=0, keepdims=True)) if d is None
                                     else np.sum(_num(a), axis=int(_num(d).ravel()[0]) - 1, keepdims=True),
            'any': lambda a: bool(np.any(_num(a) != 0)),
            'all': lambda a: bool(np.all(_num(a) != 0)),
            'double': lambda a: _num(a).astype(np.float64),
            'trace': lambda a: float(np.trace(_num(a))),
            'sort': None,                # two outputs: handled in call()
            'isfield': lambda st, name: isinstance(st, MStruct) and name in st.f,
            'reshape': self._reshape,
            'error': self._error,
            'find': self._find,
            'zeros': self._zeros,
            'eye': self._eye,
            'size': None,                # needs nargout: handled in call()
            'warning': lambda *a: np.zeros((0, 0)),
            'Inf': lambda: np.array([[np.inf]]),
            'inf': lambda: np.array([[np.inf]]),
            'true': lambda: np.array([[True]]),
            'false': lambda: np.array([[False]]),
            'pi': lambda: np.array([[np.pi]]),
        }
        self._io = self._io_builtins()
        for k, v in (stubs or {}).items():
            self.builtins[k] = v

    @staticmethod
    def _ew(f, a):
        a = _num(a)
        out = np.empty(a.shape)
        for idx in np.ndindex(a.shape):
            out[idx] = f(float(a[idx]))
        return out

    @staticmethod
    def _dims(args):
        if len(args) == 1:
            a = _num(args[0])
            if a.size == 1:
                return int(a.ravel()[0]), int(a.ravel()[0])
            return int(a.ravel()[0]), int(a.ravel()[1])
        return int(_num(args[0]).ravel()[0]), int(_num(args[1]).ravel()[0])

    def _zeros(self, *args):
        return np.zeros(self._dims(args))

    def _eye(self, *args):
        r, c = self._dims(args)
        return np.eye(r, c)

    # ---- files, paths, formatted text: what a driver script needs around the arithmetic
    @staticmethod
    def _sprintf(fmt, args):
        """MATLAB's fprintf / sprintf rule: the arguments are flattened (column-major) and the format is reused until they
        are used up; with no argument left for a conversion the output stops there; escapes are processed in the format."""
        import re
        flat = []
        for a in args:
            if isinstance(a, str):
                flat.append(a)
            else:
                flat.extend(_num(a).ravel(order='F').tolist())
        fmt = fmt.replace('\\n', '\n').replace('\\t', '\t').replace('%%', '\0')
        parts = re.split(r'(%[-+ 0#]*\d*(?:\.\d+)?[dfegs])', fmt)
        nconv = sum(1 for p in parts if p.startswith('%'))
        out, i, first = [], 0, True
        while first or i < len(flat):
            first = False
            for p in parts:
                if p.startswith('%'):
                    if i >= len(flat):
                        return ''.join(out).replace('\0', '%')
                    v = flat[i]; i += 1
                    if p.endswith('s'):
                        out.append(p % (v if isinstance(v, str) else ('%g' % v)))
                    elif p.endswith('d'):
                        out.append((p % int(v)) if float(v) == int(v) else (p[:-1] + 'e') % v)
                    else:
                        out.append(p % float(v))
                else:
                    out.append(p)
            if nconv == 0:
                break
        return ''.join(out).replace('\0', '%')

    def _io_builtins(self):
        import os

        def fopen(path, mode='r'):
            try:
                fh = open(path, mode)
            except OSError:
                return -1.0
            fid = 3 + len(self._fids)
            self._fids[fid] = fh
            return float(fid)

        def fid_of(v):
            return int(_num(v).ravel()[0])

        def fclose(fid):
            self._fids.pop(fid_of(fid)).close()

        def fprintf(*a):
            if a and isinstance(a[0], np.ndarray):
                self._fids[fid_of(a[0])].write(self._sprintf(a[1], a[2:]))
            else:
                self.printed.append(self._sprintf(a[0], a[1:]))

        def dir_(path):
            from types import SimpleNamespace
            names = ['.', '..'] + sorted(os.listdir(path))
            sa = StructArray([SimpleNamespace(name=n, isdir=os.path.isdir(os.path.join(path, n))) for n in names])
            sa.fields = ['name', 'isdir']
            return sa

        def which(name):
            if name in self.builtins or name in self._io:
                return 'built-in'
            return self._resolve(name) or ''

        def addpath(d, where=None):
            if d in self.path:
                self.path.remove(d)
            if where == '-end':
                self.path.append(d)
            else:
                self.path.insert(0, d)                      # in front, as MATLAB does

        def mfilename(kind=None):
            f = next((x for x in reversed(self._files) if x), None)
            if f is None:
                return ''
            return os.path.splitext(f)[0] if kind == 'fullpath' else os.path.splitext(os.path.basename(f))[0]

        def assert_(cond, fmt=None, *a):
            if not _truth(cond if isinstance(cond, np.ndarray) else np.array([[bool(cond)]])):
                raise MiniMatlabError('assertion failed: ' + (self._sprintf(fmt, a) if fmt else ''))

        def strfind(a, b):
            hits = [i + 1.0 for i in range(len(a) - len(b) + 1) if a[i:i + len(b)] == b] if b else []
            return np.array([hits]) if hits else np.zeros((0, 0))

        def sscanf(txt, fmt):
            if fmt != '%f':
                raise MiniMatlabError("sscanf: only '%f'")
            return np.array([float(t) for t in txt.split()]).reshape(-1, 1)

        def fread(fid, n, kind):
            if kind != 'uint8=>char':
                raise MiniMatlabError("fread: only 'uint8=>char'")
            return self._fids[fid_of(fid)].read()

        self.printed = []
        return {'fopen': fopen, 'fclose': fclose, 'fprintf': fprintf, 'fread': fread, 'sscanf': sscanf, 'dir': dir_,
                'which': which, 'addpath': addpath, 'mfilename': mfilename, 'assert': assert_, 'strfind': strfind,
                'fullfile': lambda *p: os.path.join(*p), 'fileparts': lambda p: os.path.dirname(p),
                'exist': lambda p, kind=None: 7.0 if os.path.isdir(p) else (2.0 if os.path.isfile(p) else 0.0),
                'mkdir': lambda p: os.makedirs(p, exist_ok=True), 'strtrim': lambda t: t.strip(),
                'fileread': lambda p: open(p).read(), 'strrep': lambda t, a, b: t.replace(a, b),
                'nan': lambda *a: np.full(self._dims(a) if a else (1, 1), np.nan),
                'feval': lambda name, *a: self.call(name, list(a), 1, {})[0]}

    @staticmethod
    def _struct(*args):
        if len(args) % 2:
            raise MiniMatlabError('struct: field names and values come in pairs')
        if any(isinstance(v, EmptyCell) for v in args[1::2]):           # struct('a', {}, 'b', {}): an empty struct array
            if not all(isinstance(v, EmptyCell) for v in args[1::2]):
                raise MiniMatlabError('struct: cell-array values are supported only as {}')
            sa = StructArray([])
            sa.fields = list(args[0::2])
            return sa
        return MStruct({args[i]: args[i + 1] for i in range(0, len(args), 2)})

    @staticmethod
    def _reshape(a, *dims):
        a = _num(a)
        if len(dims) == 1:
            dims = tuple(_num(dims[0]).ravel())
        d = [None if (isinstance(x, np.ndarray) and x.size == 0) else int(_num(x).ravel()[0]) for x in dims]
        if d.count(None) > 1 or len(d) != 2:
            raise MiniMatlabError('reshape: two dimensions, at most one of them []')
        if None in d:
            known = d[1 - d.index(None)]
            if known == 0 or a.size % known:
                raise MiniMatlabError('reshape: size is not divisible')
            d[d.index(None)] = a.size // known
        if d[0] * d[1] != a.size:
            raise MiniMatlabError('reshape: number of elements must not change')
        return a.reshape((d[0], d[1]), order='F').copy()

    @staticmethod
    def _error(ident, *rest):
        raise MiniMatlabError('error %s: %s' % (ident, rest[0] if rest else ''))

    # ---- classes: inheritance
    def _chain(self, cls):
        """The class and its user-defined superclasses, most derived first (`handle` ends the chain)."""
        out = [cls]
        while out[-1]['super'] in self.classes:
            out.append(self.classes[out[-1]['super']])
        return out

    def find_method(self, cls, name):
        for c in self._chain(cls):
            if name in c['methods']:
                return c['methods'][name]
        return None

    @staticmethod
    def _find(a):
        a = _num(a)
        idx = np.flatnonzero(a.ravel(order='F') != 0) + 1.0
        return idx.reshape(1, -1) if a.shape[0] <= 1 else idx.reshape(-1, 1)

    # ---- loading
    def load_source(self, src):
        node = Parser(tokenize(src)).parse_file()
        if node['kind'] == 'functions':
            self.functions.update(node['functions'])
            return sorted(node['functions'])
        if node['super'] not in (None, 'handle') and node['super'] not in self.classes and self.path:
            self._resolve(node['super'])                    # a superclass found on the search path
        node['handle'] = node['super'] == 'handle' or (node['super'] in self.classes and self.classes[node['super']]['handle'])
        self.classes[node['name']] = node
        return node['name']

    def load_file(self, path):
        with open(path) as f:
            src = f.read()
        before = set(self.functions)
        out = self.load_source(src)
        for name in (out if isinstance(out, list) else []):
            self.functions[name]['file'] = path
        if isinstance(out, str):
            for m in self.classes[out]['methods'].values():
                m['file'] = path
        return out

    def _resolve(self, name):
        """`name` on the search path (addpath): the first <dir>/<name>.m, loaded on first use as MATLAB does."""
        import os
        for d in self.path:
            f = os.path.join(d, name + '.m')
            if os.path.isfile(f):
                if f not in self._loaded:
                    self._loaded.add(f)
                    self.load_file(f)
                return f
        return None

    # ---- objects and calls
    def construct(self, name, args):
        cls = self.classes[name]
        obj = MObject(cls)
        for c in reversed(self._chain(cls)):
            for pname, default in c['props']:
                obj.props[pname] = self.eval(default, {}) if default is not None else np.zeros((0, 0))
        for c in self._chain(cls):                       # a class without a constructor hands its arguments up
            ctor = c['methods'].get(c['name'])
            if ctor is not None:
                return self.call_function(ctor, args, 1, preset={ctor['outs'][0]: obj})[0]
        return obj

    def call_function(self, f, args, nargout, preset=None):
        self._files.append(f.get('file'))
        try:
            return self._call_function(f, args, nargout, preset)
        finally:
            self._files.pop()

    def _call_function(self, f, args, nargout, preset=None):
        if len(args) > len(f['params']):
            raise MiniMatlabError('%s: too many input arguments' % f['name'])
        env = dict(preset or {})
        env['nargin'] = np.array([[float(len(args))]])
        for p, a in zip(f['params'], args):
            if p != '~':
                env[p] = _copy_value(a)
        try:
            self.exec_block(f['body'], env)
        except _Return:
            pass
        outs = []
        for o in f['outs'][:max(nargout, 1)]:
            if o not in env:
                if nargout == 0:
                    break
                raise MiniMatlabError('%s: output %s not assigned' % (f['name'], o))
            outs.append(env[o])
        return outs

    def call(self, name, args, nargout, env):
        if name == 'size':
            a = args[0]
            shp = a.shape if isinstance(a, np.ndarray) else (1, len(a.items))
            if len(args) == 2:
                return [np.array([[float(shp[int(_num(args[1]).ravel()[0]) - 1])]])]
            if nargout <= 1:
                return [np.array([[float(shp[0]), float(shp[1])]])]
            return [np.array([[float(shp[0])]]), np.array([[float(shp[1])]])]
        if name == 'sort':
            a = _num(args[0])
            if min(a.shape) > 1:
                raise MiniMatlabError('sort: vectors only')
            order = np.argsort(a.ravel(order='F'), kind='stable')
            vals = a.ravel(order='F')[order].reshape(a.shape)
            return [vals, (order + 1.0).reshape(a.shape)][:max(nargout, 1)]
        if name == 'find' and len(args) == 2:
            hit = self._find(args[0])
            n = int(_num(args[1]).ravel()[0])
            flat = hit.ravel()[:n]
            return [flat.reshape(1, -1) if hit.shape[0] <= 1 else flat.reshape(-1, 1)]
        if name == 'eig':
            # MathWorks built-in (LAPACK): a symmetric input gives ascending eigenvalues and orthonormal vectors (dsyev),
            # anything else goes through the general solver; [V,D] = eig(A) returns D as a diagonal matrix.
            a = _num(args[0])
            w, v = np.linalg.eigh(a) if np.array_equal(a, a.T) else np.linalg.eig(a)
            return [v, np.diag(w)] if nargout >= 2 else [w.reshape(-1, 1)]
        if name in self.classes:
            return [self.construct(name, args)]
        if name in self.functions:
            return self.call_function(self.functions[name], args, nargout)
        if args and isinstance(args[0], MObject) and name not in self.builtins:       # function-call form of a method: delete(obj)
            m = self.find_method(args[0].cls, name)
            if m is not None:
                return self.call_function(m, args, nargout)
        if name not in self.builtins and name not in self._io and self.path and self._resolve(name):
            if name not in self.classes and name not in self.functions:
                raise MiniMatlabError('%s.m does not define %s' % (name, name))
            return self.call(name, args, nargout, env)
        if name in self.builtins:
            fn = self.builtins[name]
            if getattr(fn, 'wants_nargout', False):        # a stand-in for a MEX file: fn(nargout, *args) -> list of outputs
                return [to_matlab(o) for o in fn(nargout, *args)]
            out = fn(*args)
            return [to_matlab(out)]
        if name in self._io:
            return [to_matlab(self._io[name](*args))]
        raise MiniMatlabError('undefined function or variable %r' % name)

    def call_value(self, fv, args, nargout):
        if isinstance(fv, BoundMethod):
            if isinstance(fv.obj, PyObj):
                out = fv.func(*[self._to_python(a) for a in args])
                return [to_matlab(out)]
            if fv.obj is None:                              # static method
                return self.call_function(fv.func, list(args), nargout)
            return self.call_function(fv.func, [fv.obj] + list(args), nargout)
        raise MiniMatlabError('value is not callable')

    @staticmethod
    def _to_python(v):
        if isinstance(v, np.ndarray):
            if v.size == 0:
                return None
            return v.ravel().copy() if min(v.shape) == 1 else v.copy()
        if isinstance(v, PyObj):
            return v.o
        return v

    # ---- statements
    def exec_block(self, stmts, env):
        for s in stmts:
            self.exec(s, env)

    def exec(self, s, env):
        k = s[0]
        try:
            if k == 'assign':
                self.assign(s[1], self.eval(s[2], env), env)
            elif k == 'massign':
                targets = s[1]
                vals = self.eval_multi(s[2], env, len(targets))
                if len(vals) < len(targets):
                    raise MiniMatlabError('not enough outputs')
                for tgt, v in zip(targets, vals):
                    self.assign(tgt, v, env)
            elif k == 'expr':
                self.eval_multi(s[1], env, 0)
            elif k == 'if':
                for cond, body in s[1]:
                    if _truth(self.eval(cond, env)):
                        self.exec_block(body, env)
                        break
                else:
                    if s[2] is not None:
                        self.exec_block(s[2], env)
            elif k == 'for':
                rng = _num(self.eval(s[2], env))
                for j in range(rng.shape[1] if rng.size else 0):
                    col = rng[:, j:j + 1]
                    env[s[1]] = col.copy() if col.shape[0] > 1 else col.reshape(1, 1).copy()
                    try:
                        self.exec_block(s[3], env)
                    except _Continue:
                        continue
                    except _Break:
                        break
            elif k == 'while':
                while _truth(self.eval(s[1], env)):
                    try:
                        self.exec_block(s[2], env)
                    except _Continue:
                        continue
                    except _Break:
                        break
            elif k == 'break':
                raise _Break()
            elif k == 'continue':
                raise _Continue()
            elif k == 'switch':
                subj = self.eval(s[1], env)
                for val, body in s[2]:
                    v = self.eval(val, env)
                    hit = (subj == v) if isinstance(subj, str) or isinstance(v, str) else bool(np.all(_num(subj) == _num(v)))
                    if hit:
                        self.exec_block(body, env)
                        break
                else:
                    if s[3] is not None:
                        self.exec_block(s[3], env)
            elif k == 'return':
                raise _Return()
            else:
                raise MiniMatlabError('statement %r' % k)
        except MiniMatlabError as e:
            if 'line' not in str(e):
                raise type(e)('line %s: %s' % (s[-1], e))
            raise

    # ---- assignment
    def assign(self, node, val, env):
        val = _copy_value(val)
        k = node[0]
        if k == 'id':
            env[node[1]] = val
        elif k == 'tilde':
            pass
        elif k == 'index' and isinstance(val, MStruct) and isinstance(self._eval_or_none(node[1], env), StructArray):
            # lm(end + 1) = struct(...): one element of a struct array
            from types import SimpleNamespace
            sa = self._eval_or_none(node[1], env)
            self._end_stack.append(len(sa.items))
            try:
                i = self._index_arg(node[2][0], env)
            finally:
                self._end_stack.pop()
            if len(node[2]) != 1 or i.size != 1 or int(i[0]) > len(sa.items):
                raise MiniMatlabError('struct-array assignment takes one scalar index, at most one past the end')
            fields = list(getattr(sa, 'fields', [])) or (sorted(vars(sa.items[0])) if sa.items else sorted(val.f))
            if sorted(fields) != sorted(val.f):
                raise MiniMatlabError('subscripted assignment between dissimilar structures')
            items = list(sa.items)
            item = SimpleNamespace(**{f: _copy_value(val.f[f]) for f in fields})
            if int(i[0]) == len(items):
                items.append(item)
            else:
                items[int(i[0])] = item
            out = StructArray(items)
            out.fields = fields
            self.assign(node[1], out, env)
        elif k == 'field' and node[1][0] == 'index' and isinstance(self._eval_or_none(node[1][1], env), StructArray):
            # lm(k).f = v on a struct array: grows it, new elements get every known field as []
            from types import SimpleNamespace
            sa = self._eval_or_none(node[1][1], env)
            i = self._index_arg(node[1][2][0], env)
            if len(node[1][2]) != 1 or i.size != 1:
                raise MiniMatlabError('struct-array assignment takes one scalar index')
            fields = list(getattr(sa, 'fields', [])) or (sorted(vars(sa.items[0])) if sa.items else [])
            if node[2] not in fields:
                fields.append(node[2])
            items = [SimpleNamespace(**{f: getattr(it, f, np.zeros((0, 0))) for f in fields}) for it in sa.items]
            while len(items) <= int(i[0]):
                items.append(SimpleNamespace(**{f: np.zeros((0, 0)) for f in fields}))
            setattr(items[int(i[0])], node[2], val)
            out = StructArray(items)
            out.fields = fields
            self.assign(node[1][1], out, env)
        elif k == 'field':
            base = self._eval_or_none(node[1], env)
            if base is None or isinstance(base, MStruct):
                base = MStruct() if base is None else _copy_value(base)
                base.f[node[2]] = val
                self.assign(node[1], base, env)
                return
            if not isinstance(base, MObject):
                raise MiniMatlabError('field assignment on a non-object')
            setter = self.find_method(base.cls, 'set.' + node[2])
            key = (id(base), 'set.' + node[2])
            if setter is not None and key not in self._in_accessor:
                self._in_accessor.add(key)
                try:
                    out = self.call_function(setter, [base, val], 1 if setter['outs'] else 0)
                finally:
                    self._in_accessor.discard(key)
                if setter['outs']:                          # value class: obj = set.p(obj, v)
                    base = out[0]
                self.assign(node[1], base, env)
                return
            if node[2] not in base.props:
                raise MiniMatlabError('no property %r' % node[2])
            base.props[node[2]] = val
            self.assign(node[1], base, env)
        elif k == 'index':
            cur = self._eval_or_none(node[1], env)
            if cur is None:
                cur = np.zeros((0, 0))
            if not isinstance(cur, np.ndarray):
                raise MiniMatlabError('indexed assignment on a non-array')
            idx = self._index_args(cur, node[2], env)
            self.assign(node[1], self._index_assign(cur, idx, val), env)
        else:
            raise MiniMatlabError('cannot assign to %r' % k)

    def _eval_or_none(self, node, env):
        if node[0] == 'id' and node[1] not in env:
            return None
        return self.eval(node, env)

    def _index_args(self, arr, nodes, env):
        """Index arguments of arr(...), with `end` standing for the last index of its position."""
        out = []
        for j, a in enumerate(nodes):
            if isinstance(arr, StructArray):
                last = len(arr.items)
            else:
                last = arr.size if len(nodes) == 1 else (arr.shape[j] if j < 2 else 1)
            self._end_stack.append(last)
            try:
                out.append(self._index_arg(a, env))
            finally:
                self._end_stack.pop()
        return out

    def _index_arg(self, a, env):
        if a[0] == 'colon_all':
            return ':'
        v = self.eval(a, env)
        if isinstance(v, np.ndarray) and v.dtype == np.bool_:
            return np.flatnonzero(v.ravel(order='F'))
        v = _num(v).ravel(order='F')
        if np.any(v != np.round(v)) or np.any(v < 1):
            raise MiniMatlabError('index must be a positive integer, got %r' % (v,))
        return v.astype(int) - 1

    @staticmethod
    def _index_assign(cur, idx, val):
        val = _num(val) if isinstance(val, np.ndarray) else val
        if not isinstance(val, np.ndarray):
            raise MiniMatlabError('only numeric indexed assignment is supported')
        cur = cur.astype(np.float64) if cur.dtype == np.bool_ else cur.copy()
        if len(idx) == 1:
            i = idx[0]
            if isinstance(i, str):
                i = np.arange(cur.size)
            need = int(i.max()) + 1 if i.size else 0
            if cur.size == 0:
                cur = np.zeros((1, need))
            elif need > cur.size:
                if cur.shape[0] == 1:
                    cur = np.hstack([cur, np.zeros((1, need - cur.shape[1]))])
                elif cur.shape[1] == 1:
                    cur = np.vstack([cur, np.zeros((need - cur.shape[0], 1))])
                else:
                    raise MiniMatlabError('linear index out of range on a matrix')
            flat = cur.ravel(order='F').copy()
            v = val.ravel(order='F')
            if v.size == 1:
                flat[i] = v[0]
            elif v.size == i.size:
                flat[i] = v
            else:
                raise MiniMatlabError('assignment size mismatch')
            return flat.reshape(cur.shape, order='F')
        if len(idx) != 2:
            raise MiniMatlabError('only 1-D / 2-D indexing')
        r, c = idx
        if isinstance(r, str):
            r = np.arange(cur.shape[0])
        if isinstance(c, str):
            c = np.arange(cur.shape[1])
        nr = max(cur.shape[0], int(r.max()) + 1 if r.size else 0)
        nc = max(cur.shape[1], int(c.max()) + 1 if c.size else 0)
        if (nr, nc) != cur.shape:
            grown = np.zeros((nr, nc))
            grown[:cur.shape[0], :cur.shape[1]] = cur
            cur = grown
        if val.size == 1:
            cur[np.ix_(r, c)] = val.ravel()[0]
        elif val.shape == (r.size, c.size):
            cur[np.ix_(r, c)] = val
        elif val.size == r.size * c.size and min(val.shape) == 1 and min(r.size, c.size) == 1:
            cur[np.ix_(r, c)] = val.reshape(r.size, c.size)
        else:
            raise MiniMatlabError('assignment size mismatch: %r into %dx%d' % (val.shape, r.size, c.size))
        return cur

    def _args(self, nodes, env):
        """Call arguments; a comma-separated list (field of a multi-element struct array) expands into several, as in MATLAB."""
        out = []
        for a in nodes:
            v = self.eval(a, env)
            if isinstance(v, CSList):
                out.extend(v.vals)
            else:
                out.append(v)
        return out

    def _super_call(self, ref, args, nargout, env):
        """`obj@Super(args)` inside a constructor runs the superclass constructor on the object under construction;
        `method@Super(obj, args)` calls the superclass's version of an overridden method."""
        _, left, sup = ref
        if sup not in self.classes:
            raise MiniMatlabError('unknown superclass %r' % sup)
        if left in env and isinstance(env[left], MObject):
            ctor = self.classes[sup]['methods'].get(sup)
            if ctor is None:
                return []
            env[left] = self.call_function(ctor, args, 1, preset={ctor['outs'][0]: env[left]})[0]
            return []
        m = self.find_method(self.classes[sup], left)
        if m is None:
            raise MiniMatlabError('%s has no method %r' % (sup, left))
        return self.call_function(m, args, nargout)

    # ---- expressions
    def eval_multi(self, node, env, nargout):
        """Evaluate a call-like expression asking for `nargout` outputs; anything else yields one value."""
        if node[0] == 'index':
            base = node[1]
            if base[0] == 'superref':
                return self._super_call(base, self._args(node[2], env), nargout, env)
            if base[0] == 'id' and base[1] not in env:
                args = self._args(node[2], env)
                return self.call(base[1], args, nargout, env)
            bv = self.eval(base, env)
            if isinstance(bv, BoundMethod):
                args = self._args(node[2], env)
                return self.call_value(bv, args, nargout)
        if node[0] == 'id' and node[1] not in env:
            return self.call(node[1], [], nargout, env)
        if node[0] == 'field':
            v = self.eval(node, env)
            if isinstance(v, BoundMethod):
                return self.call_value(v, [], nargout)
            return [v]
        return [self.eval(node, env)]

    def eval(self, node, env):
        k = node[0]
        if k == 'num':
            return np.array([[node[1]]])
        if k == 'str':
            return node[1]
        if k == 'emptycell':
            return EmptyCell()
        if k == 'end':
            if not self._end_stack:
                raise MiniMatlabError('`end` outside an index')
            return np.array([[float(self._end_stack[-1])]])
        if k == 'paren':
            v = self.eval(node[1], env)
            if isinstance(v, CSList):
                raise MiniMatlabError('comma-separated list in parentheses')
            return v
        if k == 'id':
            if node[1] in env:
                return env[node[1]]
            return self.call(node[1], [], 1, env)[0]
        if k == 'transpose':
            v = self.eval(node[1], env)
            if isinstance(v, str):                          # char arrays live here as Python strings, orientation-free
                return v
            return _num(v).T.copy()
        if k == 'un':
            a = self.eval(node[2], env)
            if node[1] == '-':
                return -_num(a)
            if node[1] == '+':
                return _num(a)
            return _num(a) == 0
        if k == 'bin':
            op = node[1]
            if op == '&&':
                return np.array([[_truth(self.eval(node[2], env)) and _truth(self.eval(node[3], env))]])
            if op == '||':
                return np.array([[_truth(self.eval(node[2], env)) or _truth(self.eval(node[3], env))]])
            a, b = self.eval(node[2], env), self.eval(node[3], env)
            return self._binop(op, a, b)
        if k == 'range':
            a = float(_num(self.eval(node[1], env)).ravel()[0])
            b = float(_num(self.eval(node[2], env)).ravel()[0])
            st = 1.0 if node[3] is None else float(_num(self.eval(node[3], env)).ravel()[0])
            n = int(np.floor((b - a) / st + 1e-10)) + 1
            return (a + st * np.arange(max(n, 0))).reshape(1, -1)
        if k == 'matrix':
            rows = []
            for r in node[1]:
                elems = []
                for e in r:
                    v = self.eval(e, env)
                    if isinstance(v, CSList):
                        elems.extend(v.vals)
                    else:
                        elems.append(v)
                elems = [_num(e) for e in elems]
                elems = [e for e in elems if e.size > 0]
                if elems:
                    rows.append(np.hstack(elems))
            if not rows:
                return np.zeros((0, 0))
            return np.vstack(rows)
        if k == 'dynfield':
            name = self.eval(node[2], env)
            if not isinstance(name, str):
                raise MiniMatlabError('dynamic field name must be a string')
            return self.eval(('field', node[1], name), env)
        if k == 'field' and node[1][0] == 'id' and node[1][1] not in env and node[1][1] in self.classes:
            m = self.find_method(self.classes[node[1][1]], node[2])                 # ClassName.method: a static method
            if m is None:
                raise MiniMatlabError('%s has no method %r' % (node[1][1], node[2]))
            return BoundMethod(None, m)
        if k == 'field':
            base = self.eval(node[1], env)
            name = node[2]
            if isinstance(base, MObject):
                getter = self.find_method(base.cls, 'get.' + name)
                key = (id(base), 'get.' + name)
                if getter is not None and key not in self._in_accessor:
                    self._in_accessor.add(key)
                    try:
                        return self.call_function(getter, [base], 1)[0]
                    finally:
                        self._in_accessor.discard(key)
                if name in base.props:
                    return base.props[name]
                m = self.find_method(base.cls, name)
                if m is not None:
                    return BoundMethod(base, m)
                raise MiniMatlabError('no property or method %r of %s' % (name, base.cls['name']))
            if isinstance(base, MStruct):
                if name not in base.f:
                    raise MiniMatlabError('reference to non-existent field %r' % name)
                return base.f[name]
            if isinstance(base, StructArray):
                vals = [to_matlab(getattr(it, name)) for it in base.items]
                return vals[0] if len(vals) == 1 else CSList(vals)
            if isinstance(base, PyObj):
                v = getattr(base.o, name)
                if callable(v):
                    return BoundMethod(base, v)
                return to_matlab(v)
            raise MiniMatlabError('field %r of a non-struct' % name)
        if k == 'index':
            base = node[1]
            if base[0] == 'id' and base[1] not in env:
                args = self._args(node[2], env)
                return self.call(base[1], args, 1, env)[0]
            bv = self.eval(base, env)
            if isinstance(bv, BoundMethod):
                args = self._args(node[2], env)
                out = self.call_value(bv, args, 1)
                if not out:
                    raise MiniMatlabError('method returned nothing')
                return out[0]
            if isinstance(bv, StructArray):
                idx = self._index_args(bv, node[2][:1], env)[0]
                return StructArray([bv.items[i] for i in idx])
            if isinstance(bv, np.ndarray):
                idx = self._index_args(bv, node[2], env)
                return self._index(bv, idx)
            if isinstance(bv, str):                         # name(1): characters of a string
                idx = self._index_arg(node[2][0], env)
                return ''.join(bv[i] for i in idx)
            raise MiniMatlabError('cannot index a %r' % type(bv))
        raise MiniMatlabError('expression %r' % k)

    @staticmethod
    def _index(a, idx):
        if len(idx) == 1:
            i = idx[0]
            flat = a.ravel(order='F')
            if isinstance(i, str):
                return flat.reshape(-1, 1).copy()
            if np.any(i >= flat.size):
                raise MiniMatlabError('index exceeds array bounds')
            out = flat[i]
            if a.shape[0] == 1 or a.size == 0:
                return out.reshape(1, -1).copy()
            if a.shape[1] == 1:
                return out.reshape(-1, 1).copy()
            return out.reshape(1, -1).copy()
        r, c = idx
        if isinstance(r, str):
            r = np.arange(a.shape[0])
        if isinstance(c, str):
            c = np.arange(a.shape[1])
        if (r.size and r.max() >= a.shape[0]) or (c.size and c.max() >= a.shape[1]):
            raise MiniMatlabError('index exceeds array bounds')
        return a[np.ix_(r, c)].copy()

    @staticmethod
    def _binop(op, a, b):
        if op in ('==', '~=') and isinstance(a, str) and isinstance(b, str):
            if len(a) != len(b) and 1 not in (len(a), len(b)):
                raise MiniMatlabError('char arrays of different lengths compared with ==')
            n = max(len(a), len(b))
            eq = np.array([[(a[i % len(a)] == b[i % len(b)]) for i in range(n)]])
            return eq if op == '==' else ~eq
        if op in ('==', '~=') and (isinstance(a, str) or isinstance(b, str)):
            raise MiniMatlabError('string comparison with == is not supported')
        a, b = _num(a), _num(b)
        if op == '+':
            return a + b
        if op == '-':
            return a - b
        if op == '.*':
            return a * b
        if op == './':
            with np.errstate(divide='ignore', invalid='ignore'):
                return a / b
        if op == '*':
            if a.size == 1 or b.size == 1:
                return a * b
            if a.shape[1] != b.shape[0]:
                raise MiniMatlabError('inner matrix dimensions must agree: %r * %r' % (a.shape, b.shape))
            return a @ b
        if op == '/':
            if b.size == 1:
                with np.errstate(divide='ignore', invalid='ignore'):      # x/0 is Inf in MATLAB, silently
                    return a / b
            raise MiniMatlabError('matrix right division is not supported')
        if op in ('^', '.^'):
            if op == '^' and a.size > 1:
                if b.size == 1 and float(b.ravel()[0]) == -1.0 and a.shape[0] == a.shape[1]:
                    return MC.inv2(a)
                raise MiniMatlabError('matrix power other than ^-1 is not supported')
            if a.size == 1 and b.size == 1:
                return np.array([[float(a.ravel()[0]) ** float(b.ravel()[0])]])
            return a ** b
        if op == '==':
            return a == b
        if op == '~=':
            return a != b
        if op == '<':
            return a < b
        if op == '<=':
            return a <= b
        if op == '>':
            return a > b
        if op == '>=':
            return a >= b
        if op == '&':
            return (a != 0) & (b != 0)
        if op == '|':
            return (a != 0) | (b != 0)
        raise MiniMatlabError('operator %r' % op)


# ------------------------------------------------------------------------------------------------------ driver
class ReferenceFilter:
    """The reference's EKF_SLAM / EKF_SLAM_UC object, executed from its own source, with the duck type the parity tests
    use (predict / measure / x / P / s).  `ref_dir` is a checkout of SamShue/EKF_SLAM; its files are read, not copied.
    RANSAC() -- called by the EKF_SLAM constructor to fill a property the filter never reads again (EKF_SLAM.m:33) -- is
    ROS-bound perception and out of scope: it is served by an empty stand-in."""

    def __init__(self, ref_dir, cls_name):
        import os
        self.curves = []                 # what plot() hands to the figure: one (xr, yr) pair per ellipse
        noop = lambda *a: None
        self.it = Interpreter(stubs={'RANSAC': lambda: np.zeros((0, 0)), 'hold': noop, 'drawRobot': noop, 'scatter': noop,
                                     'plot': lambda xr, yr: self.curves.append((_num(xr).ravel().copy(), _num(yr).ravel().copy()))})
        for f in ('Correspondence.m', 'EKF_SLAM.m', 'EKF_SLAM_UC.m'):
            self.it.load_file(os.path.join(ref_dir, f))
        self.obj = self.it.construct(cls_name, [])
        self.n_landmarks = []

    def _method(self, name, *args):
        f = self.obj.cls['methods'][name]
        return self.it.call_function(f, [self.obj] + [to_matlab(a) for a in args], 0)

    def predict(self, u):
        self._method('predict', np.asarray(u, dtype=np.float64))

    def measure(self, laserData, u, landmark_list):
        self._method('measure', laserData, np.asarray(u, dtype=np.float64), PyObj(landmark_list))
        self.n_landmarks.append((self.x.size - 3) // 2)

    def plot_curves(self):
        """Run the class's own plot(h, landmark_list) (EKF_SLAM.m:154-234) with the figure calls (hold, drawRobot, scatter,
        the landmark object's plot) as no-ops and plot(xr, yr) as a recorder: returns the covariance-ellipse polylines the
        reference would draw, robot first, as an array (1 + N, 2, len(t))."""
        from types import SimpleNamespace
        self.curves = []
        lm = SimpleNamespace(landmarkObj=SimpleNamespace(plot=lambda x, observed: None))
        self._method('plot', PyObj(lm))
        return np.array([[cx, cy] for cx, cy in self.curves])

    @property
    def x(self):
        return self.obj.props['x'].ravel().copy()

    @property
    def P(self):
        return self.obj.props['P'].copy()

    @property
    def s(self):
        return self.obj.props['s'].ravel().copy()
