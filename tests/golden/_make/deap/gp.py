"""Minimal `deap.gp` surface used by the reference grammar layer."""
from collections import defaultdict


class Primitive:
    def __init__(self, name, args, ret):
        self.name = name
        self.arity = len(args)
        self.args = args
        self.ret = ret

    def format(self, *args):
        return f"{self.name}({', '.join(args)})"


class Terminal:
    arity = 0

    def __init__(self, terminal, symbolic, ret):
        self.ret = ret
        self.value = terminal
        self.name = str(terminal)
        self.conv_fct = str if symbolic else repr

    def format(self):
        return self.conv_fct(self.value)


class PrimitiveSetTyped:
    def __init__(self, name, in_types, ret_type, prefix="ARG"):
        self.terminals = defaultdict(list)
        self.primitives = defaultdict(list)
        self.arguments = []
        self.context = {"__builtins__": None}
        self.mapping = dict()
        self.terms_count = 0
        self.prims_count = 0
        self.name = name
        self.ret = ret_type
        self.ins = in_types

    def _add(self, prim):
        raise NotImplementedError  # the reference subclass overrides this

    def addPrimitive(self, primitive, in_types, ret_type, name=None):
        if name is None:
            name = primitive.__name__
        prim = Primitive(name, in_types, ret_type)
        assert name not in self.context or self.context[name] is primitive
        self._add(prim)
        self.context[prim.name] = primitive
        self.prims_count += 1

    def addTerminal(self, terminal, ret_type, name=None):
        symbolic = False
        if name is None and callable(terminal):
            name = terminal.__name__
        assert name not in self.context
        if name is not None:
            self.context[name] = terminal
            terminal = name
            symbolic = True
        elif terminal in (True, False):
            self.context[str(terminal)] = terminal
        prim = Terminal(terminal, symbolic, ret_type)
        self._add(prim)
        self.terms_count += 1


class PrimitiveTree(list):
    def __str__(self):
        string = ""
        stack = []
        for node in self:
            stack.append((node, []))
            while len(stack[-1][1]) == stack[-1][0].arity:
                prim, args = stack.pop()
                string = prim.format(*args)
                if len(stack) == 0:
                    break
                stack[-1][1].append(string)
        return string


def compile(expr, pset):
    return eval(str(expr), pset.context, {})
