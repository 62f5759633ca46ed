"""Variables, collections, device scopes and a small deferred-execution layer.

The reference programs are TF-1 graphs: they *declare* variables and ops, then
``sess.run`` them (e.g. DOWNPOUR/DOWNPOUR.py:45-113,137).  This module gives the
same vocabulary on top of eager PyTorch:

* ``Variable`` with TF auto-naming (``Variable``, ``Variable_1`` ...), the
  ``GLOBAL_VARIABLES`` / ``LOCAL_VARIABLES`` / ``TRAINABLE_VARIABLES`` collections
  and custom collections such as ``'local_non_trainable'``
  (DOWNPOUR/DOWNPOUR.py:46-53).
* ``device(...)`` scopes, nested, with strings or device functions
  (DOWNPOUR/DOWNPOUR.py:45,86-87,202).
* ``Tensor`` / ``Op`` nodes -- thunks over torch tensors, memoised per
  ``Session.run`` -- so ``c = a + b``, ``loss = reduce_mean(square(c - target))``
  and ``sess.run([opt, c, global_step])`` read like the reference
  (Synchronous-SGD/ssgd.py:42-48,73).
* ``control_dependencies`` (DOWNPOUR/DOWNPOUR.py:67,101).

A variable's storage lives on the task its device names; every op that writes a
variable executes on that task (colocation rule,
Basics-Tutorial/Local-then-Global-Variables.ipynb:127-134,168-169).
"""
from __future__ import annotations

import contextlib
import threading
from typing import Any, Callable, Dict, Iterable, List, Optional, Sequence, Union

import torch

from .cluster import DeviceSpec, canonical_device
from . import runtime

__all__ = [
    "GraphKeys", "Graph", "get_default_graph", "reset_default_graph", "Variable",
    "Tensor", "Op", "device", "name_scope", "control_dependencies", "constant",
    "truncated_normal", "zeros", "square", "reduce_mean", "reduce_sum", "abs",
    "assign", "assign_add", "assign_sub", "no_op", "group", "op", "tensor",
    "global_variables", "local_variables", "trainable_variables", "get_collection",
    "get_collection_ref", "add_to_collection", "variables_initializer",
    "global_variables_initializer", "local_variables_initializer",
    "report_uninitialized_variables", "get_variable", "RunContext", "evaluate",
    "OpDescriptor", "stack", "set_random_seed", "OutOfRangeError", "get_global_step",
    "get_or_create_global_step", "global_step_value",
]


class GraphKeys:
    GLOBAL_VARIABLES = "variables"
    LOCAL_VARIABLES = "local_variables"
    TRAINABLE_VARIABLES = "trainable_variables"
    GLOBAL_STEP = "global_step"


class OpDescriptor:
    """What a device function sees (``replica_device_setter`` inspects ``type``)."""

    def __init__(self, type_: str, name: str, device: str = ""):
        self.type, self.name, self.device = type_, name, device


class _OpName:
    def __init__(self, name):
        self.name = name


# --------------------------------------------------------------------------- #
# Graph: names, collections, scopes
# --------------------------------------------------------------------------- #
class Graph:
    def __init__(self):
        self._collections: Dict[str, list] = {}
        self._names: Dict[str, int] = {}
        self._name_scope: List[str] = []
        self._device_stack: List[Union[str, Callable, None]] = []
        self._control_stack: List[List["Node"]] = []
        self._variables_by_name: Dict[str, "Variable"] = {}
        self.seed: Optional[int] = None
        self._seed_counter = 0

    # names ------------------------------------------------------------------ #
    def unique_name(self, base: str, mark_used: bool = True) -> str:
        scope = "/".join(self._name_scope)
        full = "%s/%s" % (scope, base) if scope else base
        n = self._names.get(full)
        if n is None:
            if mark_used:
                self._names[full] = 1
            return full
        while True:
            cand = "%s_%d" % (full, n)
            n += 1
            if cand not in self._names:
                break
        if mark_used:
            self._names[full] = n
            self._names[cand] = 1
        return cand

    # collections ------------------------------------------------------------ #
    def add_to_collection(self, key: str, value) -> None:
        self._collections.setdefault(key, []).append(value)

    def get_collection(self, key: str) -> list:
        return list(self._collections.get(key, []))

    def get_collection_ref(self, key: str) -> list:
        return self._collections.setdefault(key, [])

    # devices ---------------------------------------------------------------- #
    def resolve_device(self, opdesc: OpDescriptor) -> str:
        spec = DeviceSpec()
        for entry in self._device_stack:
            if entry is None:
                spec = DeviceSpec()
                continue
            s = entry(opdesc) if callable(entry) else entry
            spec = spec.merge_from(DeviceSpec.from_string(s or ""))
        return spec.to_string()


_default_graph = Graph()
_graph_lock = threading.Lock()


def get_default_graph() -> Graph:
    return _default_graph


def reset_default_graph() -> Graph:
    global _default_graph
    with _graph_lock:
        _default_graph = Graph()
    return _default_graph


def set_random_seed(seed: int) -> None:
    get_default_graph().seed = int(seed)


@contextlib.contextmanager
def device(spec: Union[str, Callable, None]):
    """``with device('/job:ps/task:0')`` or ``with device(replica_device_setter(...))``."""
    g = get_default_graph()
    g._device_stack.append(spec)
    try:
        yield
    finally:
        g._device_stack.pop()


@contextlib.contextmanager
def name_scope(name: str):
    g = get_default_graph()
    g._name_scope.append(name)
    try:
        yield
    finally:
        g._name_scope.pop()


@contextlib.contextmanager
def control_dependencies(deps: Sequence["Node"]):
    """Nodes created inside run ``deps`` first (once per ``Session.run``)."""
    g = get_default_graph()
    g._control_stack.append([d for d in (deps or []) if d is not None])
    try:
        yield
    finally:
        g._control_stack.pop()


# --------------------------------------------------------------------------- #
# Deferred execution
# --------------------------------------------------------------------------- #
class RunContext:
    """One ``Session.run`` call: memoises node values so that fetches listed
    together share work (one forward/backward per ``compute_gradients`` call)."""

    def __init__(self, session=None):
        self.session = session
        self.cache: Dict[int, Any] = {}
        self.feed: Dict[int, Any] = {}

    def read(self, var: "Variable") -> torch.Tensor:
        if id(var) in self.feed:
            return self.feed[id(var)]
        return var._read()


class _GradContext(RunContext):
    """Evaluates a loss with variables replaced by autograd leaves."""

    def __init__(self, parent: RunContext):
        super().__init__(parent.session)
        self.leaves: Dict[int, torch.Tensor] = {}
        self.feed = parent.feed

    def read(self, var: "Variable") -> torch.Tensor:
        leaf = self.leaves.get(id(var))
        if leaf is None:
            leaf = RunContext.read(self, var).detach().clone().float().requires_grad_(True)
            self.leaves[id(var)] = leaf
        return leaf


def evaluate(x, ctx: RunContext):
    """Value of anything fetchable under ``ctx``."""
    if isinstance(x, Node):
        return x._eval(ctx)
    if isinstance(x, Variable):
        return ctx.read(x)
    if isinstance(x, (list, tuple)):
        return type(x)(evaluate(v, ctx) for v in x)
    if isinstance(x, dict):
        return {k: evaluate(v, ctx) for k, v in x.items()}
    return x


class Node:
    def __init__(self, fn: Callable[[RunContext], Any], name: str, type_: str = "Op",
                 device_: Optional[str] = None, deps: Optional[Sequence["Node"]] = None):
        g = get_default_graph()
        self._fn = fn
        self.op = _OpName(g.unique_name(name))
        self.type = type_
        self.device = device_ if device_ is not None else canonical_or_empty(
            g.resolve_device(OpDescriptor(type_, self.op.name)))
        self._deps: List[Node] = [d for layer in g._control_stack for d in layer]
        if deps:
            self._deps.extend(deps)

    @property
    def name(self):
        return self.op.name

    def _eval(self, ctx: RunContext):
        key = id(self)
        if key in ctx.cache:
            return ctx.cache[key]
        for d in self._deps:
            d._eval(ctx)
        v = self._fn(ctx)
        ctx.cache[key] = v
        sess = ctx.session
        if sess is not None and getattr(sess, "_log_placement", False):
            sess._log_node(self)           # ConfigProto(log_device_placement=True)
        return v


def canonical_or_empty(s: str) -> str:
    return canonical_device(s) if s else ""


def _ev(x, ctx):
    if isinstance(x, (Node, Variable)):
        return evaluate(x, ctx)
    return x


class Tensor(Node):
    """Lazily evaluated value."""

    def __init__(self, fn, name="Tensor", **kw):
        super().__init__(fn, name, type_=kw.pop("type_", "Tensor"), **kw)

    def eval(self, session=None):
        return evaluate(self, RunContext(session))

    # arithmetic ------------------------------------------------------------- #
    def _bin(self, other, f, name, swap=False):
        a, b = (other, self) if swap else (self, other)
        return Tensor(lambda ctx: f(_ev(a, ctx), _ev(b, ctx)), name)

    def __add__(self, o): return self._bin(o, lambda x, y: x + y, "add")
    def __radd__(self, o): return self._bin(o, lambda x, y: x + y, "add", True)
    def __sub__(self, o): return self._bin(o, lambda x, y: x - y, "sub")
    def __rsub__(self, o): return self._bin(o, lambda x, y: x - y, "sub", True)
    def __mul__(self, o): return self._bin(o, lambda x, y: x * y, "mul")
    def __rmul__(self, o): return self._bin(o, lambda x, y: x * y, "mul", True)
    def __truediv__(self, o): return self._bin(o, lambda x, y: x / y, "truediv")
    def __rtruediv__(self, o): return self._bin(o, lambda x, y: x / y, "truediv", True)
    def __neg__(self): return Tensor(lambda ctx: -_ev(self, ctx), "Neg")
    def __getitem__(self, i): return Tensor(lambda ctx: _ev(self, ctx)[i], "strided_slice")
    __hash__ = Node.__hash__


class Op(Node):
    """Side-effecting node; evaluates to ``None`` (like fetching a TF op)."""

    def __init__(self, fn, name="op", **kw):
        def run(ctx, _fn=fn):
            _fn(ctx)
            return None
        super().__init__(run, name, type_=kw.pop("type_", "NoOp"), **kw)

    def run(self, session=None):
        return evaluate(self, RunContext(session))


def op(fn: Callable[[], Any], name: str = "py_op") -> Op:
    """Wrap a plain Python callable as a runnable op."""
    return Op(lambda ctx: fn(), name)


def tensor(fn: Callable[[], Any], name: str = "py_tensor") -> Tensor:
    """Wrap a plain Python callable as a fetchable value."""
    return Tensor(lambda ctx: fn(), name)


def no_op(name="NoOp") -> Op:
    return Op(lambda ctx: None, name)


def group(*nodes, name="group_deps") -> Op:
    flat = []
    for n in nodes:
        flat.extend(n if isinstance(n, (list, tuple)) else [n])
    return Op(lambda ctx: [evaluate(n, ctx) for n in flat] and None, name)


# --------------------------------------------------------------------------- #
# Value constructors / math
# --------------------------------------------------------------------------- #
def _next_generator() -> Optional[torch.Generator]:
    g = get_default_graph()
    if g.seed is None:
        return None
    gen = torch.Generator()
    gen.manual_seed(g.seed * 1000003 + g._seed_counter)
    g._seed_counter += 1
    return gen


def constant(value, shape=None, dtype=torch.float32) -> torch.Tensor:
    """``tf.constant(0., shape=[2])`` (DOWNPOUR/DOWNPOUR.py:46)."""
    t = torch.as_tensor(value, dtype=dtype)
    if shape is not None:
        t = t.expand(tuple(shape)).clone() if t.dim() == 0 else t.reshape(tuple(shape))
    return t


def zeros(shape, dtype=torch.float32) -> torch.Tensor:
    return torch.zeros(tuple(shape), dtype=dtype)


def truncated_normal(shape, mean=0.0, stddev=1.0, dtype=torch.float32) -> torch.Tensor:
    """Normal truncated at +-2 sigma (Non-Distributed_Setup.py:13-14)."""
    t = torch.empty(tuple(shape), dtype=torch.float32)
    torch.nn.init.trunc_normal_(t, mean=mean, std=stddev, a=mean - 2 * stddev,
                                b=mean + 2 * stddev, generator=_next_generator())
    return t.to(dtype)


def _unary(x, f, name):
    if isinstance(x, (Node, Variable)):
        return Tensor(lambda ctx: f(_ev(x, ctx)), name)
    return f(x)


def square(x): return _unary(x, lambda v: v * v, "Square")
def abs(x): return _unary(x, lambda v: torch.abs(v), "Abs")  # noqa: A001
def reduce_mean(x, axis=None):
    return _unary(x, lambda v: _reduce(v, axis, torch.mean), "Mean")
def reduce_sum(x, axis=None):
    return _unary(x, lambda v: _reduce(v, axis, torch.sum), "Sum")


def _reduce(v, axis, f):
    if isinstance(v, (list, tuple)):
        v = stack_values(v)
    return f(v) if axis is None else f(v, dim=axis)


def stack_values(vals):
    if len(vals) and isinstance(vals[0], (list, tuple)):
        return torch.stack([torch.stack(list(v)) for v in vals])
    return torch.stack(list(vals))


def stack(xs, name="stack"):
    """Lazy ``tf.stack`` over (possibly nested) lists of nodes, as used by
    ``tf.reduce_sum(grad_list, axis=0)`` (DOWNPOUR/DOWNPOUR.py:77)."""
    return Tensor(lambda ctx: stack_values(evaluate(xs, ctx)), name)


# --------------------------------------------------------------------------- #
# Variable
# --------------------------------------------------------------------------- #
class Variable:
    """Named, placed, collection-tracked tensor.

    Default collections follow TF: ``[GLOBAL_VARIABLES]`` plus
    ``TRAINABLE_VARIABLES`` when ``trainable``.  A variable created with
    ``collections=[LOCAL_VARIABLES]`` is per-worker: not shared, not saved,
    initialised by the *local* init op (DOWNPOUR/DOWNPOUR.py:46-49).
    """

    def __init__(self, initial_value=None, trainable: bool = True,
                 collections: Optional[Sequence[str]] = None, name: Optional[str] = None,
                 dtype: Optional[torch.dtype] = None, shape=None):
        g = get_default_graph()
        self._graph = g
        if callable(initial_value) and not isinstance(initial_value, torch.Tensor):
            self._init_fn = initial_value
            probe = initial_value()
        else:
            probe = torch.as_tensor(initial_value)
            self._init_fn = None
        if dtype is not None:
            probe = probe.to(dtype)
        elif probe.dtype == torch.float64:
            probe = probe.float()
        if shape is not None:
            probe = probe.reshape(tuple(shape))
        self._initial_value = probe
        self.dtype = probe.dtype
        self.shape = tuple(probe.shape)
        self.trainable = bool(trainable)
        self.op = _OpName(g.unique_name(name or "Variable"))
        if collections is None:
            collections = [GraphKeys.GLOBAL_VARIABLES]
        collections = list(collections)
        if trainable and GraphKeys.TRAINABLE_VARIABLES not in collections:
            collections.append(GraphKeys.TRAINABLE_VARIABLES)
        self.collections = collections
        for c in collections:
            g.add_to_collection(c, self)
        g._variables_by_name[self.op.name] = self
        self._device_spec = self._place(g, collections)
        self.initializer = Op(lambda ctx: self._assign_raw(self.initial_value()),
                              self.op.name + "/Assign", device_=self.device, type_="Assign")

    # placement -------------------------------------------------------------- #
    def _place(self, g: Graph, collections) -> DeviceSpec:
        spec = DeviceSpec.from_string(g.resolve_device(OpDescriptor("VariableV2", self.op.name)))
        srv = runtime.current_server()
        if spec.job is None:
            # Shared variables with no job in their spec go to the parameter
            # server (Hogwild places vars with '/cpu:0' only and
            # relies on the placer choosing /job:ps); per-worker ones stay here.
            shared = GraphKeys.GLOBAL_VARIABLES in collections
            cluster = getattr(srv, "cluster", None)
            if shared and cluster is not None and "ps" in cluster.jobs:
                spec.job, spec.task = "ps", spec.task if spec.task is not None else 0
            else:
                spec.job, spec.task = srv.job_name, srv.task_index
        if spec.task is None:
            spec.task = 0
        return spec

    @property
    def device(self) -> str:
        return canonical_device(self._device_spec)

    @property
    def name(self) -> str:
        return self.op.name + ":0"

    def _host(self):
        return runtime.current_server().host_for(self._device_spec.job, self._device_spec.task)

    @property
    def storage_key(self) -> str:
        """Key on the hosting task.  Per-worker (non-shared) variables are
        private to their process, so the plain name is unambiguous there."""
        return self.op.name

    # value access ----------------------------------------------------------- #
    def initial_value(self) -> torch.Tensor:
        if self._init_fn is not None:
            v = self._init_fn()
            return v.to(self.dtype).reshape(self.shape)
        return self._initial_value

    def _read(self) -> torch.Tensor:
        return self._host().read(self.storage_key)

    def _assign_raw(self, value) -> None:
        v = torch.as_tensor(value).to(self.dtype).reshape(self.shape)
        self._host().assign(self.storage_key, v)

    def read_value(self) -> torch.Tensor:
        return self._read()

    value = read_value

    def eval(self, session=None) -> torch.Tensor:
        return self._read()

    def is_initialized(self) -> bool:
        return not self._host().uninitialized([self.storage_key])

    def assign(self, value) -> "Variable":
        self._assign_raw(value)
        return self

    def assign_add(self, delta) -> "Variable":
        self._host().assign_add(self.storage_key, torch.as_tensor(delta).to(self.dtype))
        return self

    def assign_sub(self, delta) -> "Variable":
        self._host().assign_add(self.storage_key, -torch.as_tensor(delta).to(self.dtype))
        return self

    # lazy arithmetic (so ``c = a + b`` builds a Tensor) --------------------- #
    def _t(self) -> Tensor:
        return Tensor(lambda ctx: ctx.read(self), self.op.name + "/read", device_=self.device)

    def __add__(self, o): return self._t() + o
    def __radd__(self, o): return o + self._t()
    def __sub__(self, o): return self._t() - o
    def __rsub__(self, o): return o - self._t()
    def __mul__(self, o): return self._t() * o
    def __rmul__(self, o): return o * self._t()
    def __truediv__(self, o): return self._t() / o
    def __neg__(self): return -self._t()

    def __repr__(self):
        return "<Variable '%s' shape=%s dtype=%s device=%s>" % (
            self.name, self.shape, str(self.dtype).replace("torch.", ""), self.device)


def get_variable(name: str, shape=None, dtype=torch.float32, initializer=None,
                 trainable: bool = True, collections=None) -> Variable:
    """``tf.get_variable('g/'+v.op.name, ...)`` (DOWNPOUR/DOWNPOUR.py:205-210):
    returns the existing variable of that name or creates it (zeros unless an
    initializer is given)."""
    g = get_default_graph()
    scope = "/".join(g._name_scope)
    full = "%s/%s" % (scope, name) if scope else name
    if full in g._variables_by_name:
        return g._variables_by_name[full]
    if initializer is None:
        init = torch.zeros(tuple(shape or ()), dtype=dtype)
    elif callable(initializer):
        init = lambda: initializer(tuple(shape or ())).to(dtype)  # noqa: E731
    else:
        init = torch.as_tensor(initializer, dtype=dtype)
    return Variable(init, trainable=trainable, collections=collections, name=name, dtype=dtype)


# --------------------------------------------------------------------------- #
# Collection helpers / init ops
# --------------------------------------------------------------------------- #
def get_collection(key: str) -> list:
    return get_default_graph().get_collection(key)


def get_collection_ref(key: str) -> list:
    return get_default_graph().get_collection_ref(key)


def add_to_collection(key: str, value) -> None:
    get_default_graph().add_to_collection(key, value)


def global_variables() -> List[Variable]:
    return get_collection(GraphKeys.GLOBAL_VARIABLES)


def local_variables() -> List[Variable]:
    return get_collection(GraphKeys.LOCAL_VARIABLES)


def trainable_variables() -> List[Variable]:
    return get_collection(GraphKeys.TRAINABLE_VARIABLES)


def variables_initializer(var_list: Iterable[Variable], name="init") -> Op:
    vs = list(var_list)
    return Op(lambda ctx: [v._assign_raw(v.initial_value()) for v in vs] and None, name)


def global_variables_initializer() -> Op:
    g = get_default_graph()
    # collection is read at run time, like TF reads the graph at build time of
    # the *last* call -- scripts call this after all variables exist.
    return Op(lambda ctx: [v._assign_raw(v.initial_value())
                           for v in g.get_collection(GraphKeys.GLOBAL_VARIABLES)] and None,
              "init")


def local_variables_initializer() -> Op:
    g = get_default_graph()
    return Op(lambda ctx: [v._assign_raw(v.initial_value())
                           for v in g.get_collection(GraphKeys.LOCAL_VARIABLES)] and None,
              "init_local")


def report_uninitialized_variables(var_list: Optional[Iterable[Variable]] = None) -> Tensor:
    """[TF] the ``ready_op``: names of global variables that are not initialised."""
    g = get_default_graph()

    def fn(ctx):
        vs = list(var_list) if var_list is not None else g.get_collection(GraphKeys.GLOBAL_VARIABLES)
        by_host: Dict[Any, List[Variable]] = {}
        for v in vs:
            by_host.setdefault((v._device_spec.job, v._device_spec.task), []).append(v)
        missing = []
        for _, group_ in by_host.items():
            missing += group_[0]._host().uninitialized([v.storage_key for v in group_])
        return missing
    return Tensor(fn, "report_uninitialized_variables")


# --------------------------------------------------------------------------- #
# Assign ops (deferred forms)
# --------------------------------------------------------------------------- #
def assign(ref: Variable, value, name="Assign") -> Tensor:
    """Deferred ``tf.assign``; colocated with ``ref``."""
    def fn(ctx):
        v = _ev(value, ctx)
        ref._assign_raw(v.detach() if isinstance(v, torch.Tensor) else v)
        return ref._read()
    return Tensor(fn, name, device_=ref.device, type_="Assign")


def assign_add(ref: Variable, value, name="AssignAdd") -> Tensor:
    def fn(ctx):
        ref.assign_add(_ev(value, ctx))
        return ref._read()
    return Tensor(fn, name, device_=ref.device, type_="AssignAdd")


def assign_sub(ref: Variable, value, name="AssignSub") -> Tensor:
    def fn(ctx):
        ref.assign_sub(_ev(value, ctx))
        return ref._read()
    return Tensor(fn, name, device_=ref.device, type_="AssignSub")


# --------------------------------------------------------------------------- #
# global_step lookup / errors
# --------------------------------------------------------------------------- #
class OutOfRangeError(RuntimeError):
    """Token queue closed: the chief finished ([TF] ends the session cleanly)."""


def get_global_step() -> Optional[Variable]:
    """[TF] looks the step tensor up by the name ``global_step`` -- the scripts
    create it as a plain ``tf.Variable(..., name='global_step')``
    (DOWNPOUR/DOWNPOUR.py:88)."""
    g = get_default_graph()
    col = g.get_collection(GraphKeys.GLOBAL_STEP)
    if col:
        return col[0]
    return g._variables_by_name.get("global_step")


def get_or_create_global_step() -> Variable:
    v = get_global_step()
    if v is None:
        v = Variable(torch.zeros((), dtype=torch.int64), trainable=False, name="global_step")
        add_to_collection(GraphKeys.GLOBAL_STEP, v)
    return v


def global_step_value(v: Optional[Variable] = None) -> int:
    v = v or get_global_step()
    return int(v.read_value().item())


