"""Parse the ExaSlang text the reference emits (exastencils.py:684-925) into a
normalised statement sequence, so that it can be compared with `lowering.lower()`."""
import re

_SOLVE = re.compile(r"solve locally at (\S+)@(\d+) (with jacobi )?relax (\S+) \{")
_UNKNOWN = re.compile(r"^(\S+?)@(\d+)@\[([^\]]*)\] =>")
_PROLONG = re.compile(r"^(\S+)@(\d+) \+= (\S+) \* \((\S+)@(\d+) \* (\S+)@(\d+)\)$")
_RESTRICT = re.compile(r"^(\S+)@(\d+) = (\S*[Rr]estriction\S*)@(\d+) \* (\S+)@(\d+)$")
_RESIDUAL = re.compile(r"^gen_residual_(\S+)@(\d+) = (\S+)@(\d+) - ")
_ZERO = re.compile(r"^(\S+)@(\d+) = 0$")
_CALL = re.compile(r"^gen_mgCycle@(\d+)\(\)$")
_COPY = re.compile(r"^(\S+)@(\d+) = (\S+)@(\d+)$")


def field_index(name, field_names):
    base = name[len("gen_error_"):] if name.startswith("gen_error_") else name
    return field_names.index(base)


def parse_listing(text, field_names):
    """-> list of tuples, consecutive per-field statements of one kind merged."""
    lines = [l.strip() for l in text.splitlines()]
    assert lines[0].startswith("Function gen_mgCycle@")
    out = []
    i = 1
    colored = False
    pending_cgs = None
    while i < len(lines):
        line = lines[i]
        i += 1
        if not line or line == "}":
            continue
        if line == "color with {":
            colored = True
            i += 1  # colour expression line
            continue
        m = _SOLVE.match(line)
        if m:
            level = int(m.group(2))
            jacobi = m.group(3) is not None
            omega = float(m.group(4))
            unknowns = []
            while lines[i] != "}":
                u = _UNKNOWN.match(lines[i])
                assert u, lines[i]
                idx = tuple(int(t) for t in u.group(3).split(","))
                unknowns.append((field_index(u.group(1), field_names), idx))
                i += 1
            i += 1  # closing brace of solve locally
            if colored:
                assert lines[i] == "}"
                i += 1
            out.append(("smooth", level, omega, jacobi, colored, tuple(unknowns)))
            colored = False
            continue
        m = _PROLONG.match(line)
        if m:
            item = ("prolong_add", int(m.group(2)), float(m.group(3)))
            if not out or out[-1] != item:
                out.append(item)
            continue
        m = _RESIDUAL.match(line)
        if m:
            item = ("residual", int(m.group(2)))
            if not out or out[-1] != item:
                out.append(item)
            continue
        m = _RESTRICT.match(line)
        if m:
            item = ("restrict", int(m.group(2)))
            if not out or out[-1] != item:
                out.append(item)
            continue
        m = _CALL.match(line)
        if m:
            out.append(("cgs", int(m.group(1))))
            pending_cgs = int(m.group(1))
            continue
        m = _ZERO.match(line)
        if m:
            level = int(m.group(2))
            # `gen_error_f@min = 0` right before the coarse solver call belongs to it
            j = i
            belongs_to_cgs = False
            while j < len(lines):
                if _CALL.match(lines[j]):
                    belongs_to_cgs = int(_CALL.match(lines[j]).group(1)) == level
                    break
                if not (_ZERO.match(lines[j]) or lines[j].startswith("gen_rhs_")):
                    break
                j += 1
            if belongs_to_cgs:
                continue
            item = ("zero", level)
            if not out or out[-1] != item:
                out.append(item)
            continue
        m = _COPY.match(line)
        if m:
            # gen_rhs_f@min = RHS_f@min (before the call) / gen_error_f@min = gen_error_f@min (after)
            continue
        raise AssertionError(f"unparsed line: {line}")
    return out


def parse_equation_coefficients(text, field_names):
    """First occurrence of each (level, field) local equation -> {(field_j, offset): value}.
    Only point equations `[0, 0(, 0)]` are collected (they print the full operator row)."""
    out = {}
    for raw in text.splitlines():
        line = raw.strip()
        u = _UNKNOWN.match(line)
        if not u:
            continue
        idx = tuple(int(t) for t in u.group(3).split(","))
        if any(idx):
            continue
        level = int(u.group(2))
        fi = field_index(u.group(1), field_names)
        if (level, fi) in out:
            continue
        body = line[line.index("=> (") + 4: line.rindex(") ==")]
        coeffs = {}
        # split at top-level " + " / " - "
        tokens, depth, cur = [], 0, ""
        for k, ch in enumerate(body):
            if ch == "(":
                depth += 1
            elif ch == ")":
                depth -= 1
            if depth == 0 and ch in "+-" and k > 0 and body[k - 1] == " " and body[k + 1] == " ":
                tokens.append(cur)
                cur = ch
            else:
                cur += ch
        tokens.append(cur)
        for tok in tokens:
            tok = tok.strip()
            sign = 1.0
            if tok.startswith("+ "):
                tok = tok[2:]
            elif tok.startswith("- "):
                sign = -1.0
                tok = tok[2:]
            elif tok.startswith("-"):
                sign = -1.0
                tok = tok[1:]
            m = re.match(r"^(?:(.+?)\*)?([A-Za-z_]\w*)@(\d+)@\[([^\]]*)\](?:\*(.+))?$", tok)
            assert m, tok
            val = m.group(1) or m.group(5) or "1.0"
            val = complex(val.strip("()").replace(" ", "").replace("*I", "j").replace("I", "1j"))
            val = val.real if val.imag == 0 else val
            off = tuple(int(t) for t in m.group(4).split(","))
            coeffs[(field_index(m.group(2), field_names), off)] = sign * val
        out[(level, fi)] = coeffs
    return out
