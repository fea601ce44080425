"""One-pass lowering of a whole `batch_assign` sweep to task records.

The reference compiles a sweep task by task (`ir/routine/bloqade.py:184-202` `_generate_ir`: assign ->
AssignToLiteral -> Canonicalizer -> `EmulatorProgramCodeGen`; ~8 ms of Python per task, then the per-task
operator build).  Once a task takes ~1 ms on the GPU that loop is the bottleneck, so the B200 routine lowers
the sweep ONCE: the flattened, still symbolic circuit is walked a single time with per-task VECTOR numbers
(`vecnum.DecVec`: one Decimal per task), through

* `eval_scalar`      - the scalar expression semantics of `ir/scalar.py:284-580`,
* `convert_waveform` - the waveform nodes of `ir/control/waveform.py:255-800` -> `bloqade_analog_b200.ir`,
* `canonicalize`     - the waveform / field rewrite rules of `compiler/rewrite/common/canonicalize.py:205-372`
                       restated on VALUES (after assignment every scalar is a literal, so the reference's
                       structural tests `is_zero / is_one / ==` are value tests),
* `build_program`    - the term construction of `compiler/codegen/python/emulator_ir.py:77-262,330-382`,

and then through the SAME `flatten_waveform` / `pack_task_record` the task-by-task path uses, which yields
records[n_tasks, L] whose row k is bit-identical to the record of task k compiled by the reference pipeline.
Every branch taken must be the same for all tasks; otherwise (or for node types that need per-task work:
PythonFn, Smooth, Sample, Record, run-time vectors that vary) `NonUniformSweep` is raised and the routine
falls back to the task-by-task path.  The routine also cross-checks sampled tasks of every fast-lowered sweep
against the reference pipeline bit for bit (`routine._lower`)."""
from __future__ import annotations

from decimal import Decimal
from typing import Any, Dict, List, Optional, Sequence, Tuple

from . import ir
from .vecnum import DecVec, NonUniformSweep, dec_vector, task_value, tofloat, vmax


# ---------------------------------------------------------------------------------------------------------
# scalars (ir/scalar.py)
# ---------------------------------------------------------------------------------------------------------
def eval_scalar(expr, env: Dict[str, Any]):
    """value of a reference scalar expression for every task: Decimal (same for all tasks) or DecVec"""
    import bloqade.analog.ir.scalar as S

    if isinstance(expr, S.Literal):
        return expr.value
    if isinstance(expr, S.AssignedVariable):
        return expr.value
    if isinstance(expr, S.Variable):  # scalar.py:311-315
        if expr.name not in env:
            raise ValueError(f"Variable {expr.name} not assigned")
        value = env[expr.name]
        if isinstance(value, (list, tuple)):
            raise NonUniformSweep("vector-valued variable used as a scalar")
        return value
    if isinstance(expr, S.Negative):
        return -eval_scalar(expr.expr, env)
    if isinstance(expr, S.Add):
        return eval_scalar(expr.lhs, env) + eval_scalar(expr.rhs, env)
    if isinstance(expr, S.Mul):
        return eval_scalar(expr.lhs, env) * eval_scalar(expr.rhs, env)
    if isinstance(expr, S.Div):
        return eval_scalar(expr.lhs, env) / eval_scalar(expr.rhs, env)
    if isinstance(expr, S.Min):
        return min(eval_scalar(e, env) for e in expr.exprs)
    if isinstance(expr, S.Max):
        return max(eval_scalar(e, env) for e in expr.exprs)
    if isinstance(expr, S.Slice):  # scalar.py:455-497
        dur = eval_scalar(expr.expr, env)
        start = eval_scalar(expr.interval.start, env) if expr.interval.start is not None else Decimal("0")
        stop = eval_scalar(expr.interval.stop, env) if expr.interval.stop is not None else dur
        if start < 0:
            raise ValueError(f"Slice start must be non-negative, got {start}")
        if stop > dur:
            raise ValueError(f"Slice stop must be smaller or equal to than duration {dur}, got {stop}")
        ret = stop - start
        if ret < 0:
            raise ValueError(f"start is larger than stop, get start = {start} and stop = {stop}")
        return ret
    raise NonUniformSweep(f"scalar node {type(expr).__name__}")


# ---------------------------------------------------------------------------------------------------------
# waveforms (ir/control/waveform.py) -> bloqade_analog_b200.ir, numbers per task
# ---------------------------------------------------------------------------------------------------------
def convert_waveform(node, env) -> ir.Waveform:
    import bloqade.analog.ir.control.waveform as W

    ev = lambda e: eval_scalar(e, env)  # noqa: E731
    if isinstance(node, W.Linear):
        return ir.Linear(ev(node.start), ev(node.stop), ev(node.duration))
    if isinstance(node, W.Constant):
        return ir.Constant(ev(node.value), ev(node.duration))
    if isinstance(node, W.Poly):
        return ir.Poly([ev(c) for c in node.coeffs], ev(node.duration))
    if isinstance(node, W.Append):
        return ir.Append([convert_waveform(w, env) for w in node.waveforms])
    if isinstance(node, W.Slice):
        inner = convert_waveform(node.waveform, env)
        start = ev(node.interval.start) if node.interval.start is not None else None
        stop = ev(node.interval.stop) if node.interval.stop is not None else None
        s0 = Decimal("0") if start is None else start
        s1 = inner.duration if stop is None else stop
        if s0 < 0 or s1 > inner.duration or s1 - s0 < 0:
            raise ValueError("invalid waveform slice")
        return ir.Slice(inner, start, stop)
    if isinstance(node, W.Negative):
        return ir.Negative(convert_waveform(node.waveform, env))
    if isinstance(node, W.Scale):
        return ir.Scale(ev(node.scalar), convert_waveform(node.waveform, env))
    if isinstance(node, W.Add):
        return ir.Add(convert_waveform(node.left, env), convert_waveform(node.right, env))
    # PythonFn, Smooth, Sample (tabulated / sampled per task), Record (assigns a variable), AlignedWaveform
    raise NonUniformSweep(f"waveform node {type(node).__name__} is lowered task by task")


def _is_zero(x) -> bool:
    return x == 0


def canonicalize(node: ir.Waveform) -> ir.Waveform:
    """`Canonicalizer` waveform rules (canonicalize.py:205-331) on assigned values"""
    if isinstance(node, ir.Append):
        first_pass: List[ir.Waveform] = []
        for sub in map(canonicalize, node.waveforms):
            if _is_zero(sub.duration):
                continue
            if isinstance(sub, ir.Append):
                first_pass.extend(sub.waveforms)
            else:
                first_pass.append(sub)
        merged = [first_pass[0]]
        for nxt in first_pass[1:]:
            last = merged.pop()
            if isinstance(last, ir.Constant) and isinstance(nxt, ir.Constant) and last.value == nxt.value:
                merged.append(ir.Constant(last.value, last.duration + nxt.duration))
            else:
                merged.append(last)
                merged.append(nxt)
        return merged[0] if len(merged) == 1 else ir.Append(merged)
    if isinstance(node, ir.Add):
        left, right = canonicalize(node.left), canonicalize(node.right)
        l_const, r_const = isinstance(left, ir.Constant), isinstance(right, ir.Constant)
        l_scale, r_scale = isinstance(left, ir.Scale), isinstance(right, ir.Scale)
        if left == right:
            return canonicalize(ir.Scale(2, left))
        if _is_zero(left.duration) or (l_const and _is_zero(left.value)):
            return right
        if _is_zero(right.duration) or (r_const and _is_zero(right.value)):
            return left
        if l_const and r_const and left.duration == right.duration:
            return ir.Constant(left.value + right.value, left.duration)
        if l_scale and r_scale and left.scalar == right.scalar:
            return canonicalize(ir.Scale(left.scalar, canonicalize(ir.Add(left.waveform, right.waveform))))
        if l_scale and r_scale and left.waveform == right.waveform:
            return canonicalize(ir.Scale(left.scalar + right.scalar, left.waveform))
        if l_scale and left.waveform == right:
            return canonicalize(ir.Scale(left.scalar + 1, left.waveform))
        if r_scale and right.waveform == left:
            return canonicalize(ir.Scale(right.scalar + 1, right.waveform))
        return ir.Add(left, right)
    if isinstance(node, ir.Scale):
        sub = canonicalize(node.waveform)
        if _is_zero(node.scalar):
            return ir.Constant(0, sub.duration)
        if node.scalar == 1:
            return sub
        if isinstance(sub, ir.Scale):
            return ir.Scale(node.scalar * sub.scalar, sub.waveform)
        if isinstance(sub, ir.Constant):
            return ir.Constant(node.scalar * sub.value, sub.duration)
        return ir.Scale(node.scalar, sub)
    if isinstance(node, ir.Slice):
        sub = canonicalize(node.waveform)
        if getattr(node, "explicit_start", True) and _is_zero(node.start) and node.stop == sub.duration:
            return sub
        if isinstance(sub, ir.Scale):
            return ir.Scale(sub.scalar, _slice_like(node, sub.waveform))
        if isinstance(sub, ir.Negative):
            return ir.Negative(_slice_like(node, sub.waveform))
        return _slice_like(node, sub)
    if isinstance(node, ir.Negative):
        sub = canonicalize(node.waveform)
        if isinstance(sub, ir.Negative):
            return sub.waveform
        if isinstance(sub, ir.Constant):
            return ir.Constant(-sub.value, sub.duration)
        return ir.Negative(sub)
    return node


def _slice_like(node: ir.Slice, inner: ir.Waveform) -> ir.Slice:
    return ir.Slice(inner, node.start, node.stop)


# ---------------------------------------------------------------------------------------------------------
# fields, pulses, register (emulator_ir.py)
# ---------------------------------------------------------------------------------------------------------
class _Drives:
    """the drives of one field after assignment + Canonicalizer.visit_field_Field (canonicalize.py:337-372):
    waveforms canonicalised, spatial modulations that share a waveform merged (scaled locations summed)"""

    def __init__(self, field_node, env):
        import bloqade.analog.ir.control.field as F

        # grouped by waveform EQUALITY, compared pairwise (not hashed): waveforms that coincide for some tasks of
        # the sweep only must raise NonUniformSweep rather than be told apart by their hash
        inverse: List[Tuple[ir.Waveform, List[Any]]] = []
        for sm, wf in field_node.drives.items():
            cwf = canonicalize(convert_waveform(wf, env))
            for other, sms in inverse:
                if other == cwf:
                    sms.append(sm)
                    break
            else:
                inverse.append((cwf, [sm]))
        self.items: List[Tuple[Any, ir.Waveform]] = []
        for cwf, sms in inverse:
            others, scaled = [], {}
            for sm in sms:
                if isinstance(sm, F.ScaledLocations):
                    for loc, scale in sm.value.items():
                        scaled[loc.value] = scaled.get(loc.value, Decimal("0")) + eval_scalar(scale, env)
                else:
                    others.append(sm)
            for sm in others:
                self.items.append((sm, cwf))
            if scaled:
                self.items.append((("scaled", scaled), cwf))

    def __len__(self):
        return len(self.items)


class _Builder:
    """restatement of EmulatorProgramCodeGen (compiler/codegen/python/emulator_ir.py:33-397) on vector numbers"""

    def __init__(self, circuit, env, blockade_radius, use_hyperfine):
        import bloqade.analog.ir.control.sequence as Q
        from bloqade.analog.ir.location.location import SiteFilling

        self.env = env
        reg = circuit.register
        if not hasattr(reg, "enumerate"):
            raise NonUniformSweep("register type")
        self.sites, self.filling, self.original_index, self.positions = [], [], [], []
        for org_index, info in enumerate(reg.enumerate()):
            filled = info.filling == SiteFilling.filled
            pos = tuple(eval_scalar(p, env) for p in info.position)
            self.filling.append(filled)
            self.sites.append(pos)
            if filled:
                self.positions.append(pos)
                self.original_index.append(org_index)
        self.n_atoms = len(self.positions)
        self.n_sites = len(self.sites)
        seq = circuit.sequence
        if type(seq) is not Q.Sequence:
            raise NonUniformSweep("sequence is not flat")
        self.is_hyperfine = bool(use_hyperfine) or (Q.hyperfine in seq.pulses)  # is_hyperfine.py:8-23
        self.duration = 0.0
        self.pulses: Dict[ir.LevelCoupling, ir.Fields] = {}
        mapping = {Q.hyperfine: ir.LevelCoupling.Hyperfine, Q.rydberg: ir.LevelCoupling.Rydberg}
        for level_coupling, sub_pulse in seq.pulses.items():
            self.pulses[mapping[level_coupling]] = self.visit_pulse(sub_pulse)

    # -- spatial modulations (emulator_ir.py:330-382) ------------------------------------------------------
    def target_atoms(self, sm) -> Dict[int, Any]:
        import bloqade.analog.ir.control.field as F

        if isinstance(sm, tuple):  # merged scaled locations
            values = sm[1]
            for location in values:
                if location >= self.n_sites or location < 0:
                    raise ValueError(f"Location {location} is out of bounds for register with {self.n_sites} sites.")
            return {new: values[org] for new, org in enumerate(self.original_index) if org in values}
        if isinstance(sm, F.UniformModulation):
            return {atom: Decimal("1.0") for atom in range(self.n_atoms)}
        if isinstance(sm, (F.RunTimeVector, F.AssignedRunTimeVector)):
            value = sm.value if isinstance(sm, F.AssignedRunTimeVector) else self.env.get(sm.name)
            if value is None:
                raise ValueError(f"Variable {sm.name} not assigned")
            if not isinstance(value, (list, tuple)) or any(isinstance(v, DecVec) for v in value):
                raise NonUniformSweep("run-time vector varies over the sweep")
            for org in self.original_index:
                if org >= len(value):
                    raise ValueError(f"Index {org} is out of bounds for the runtime vector {sm.name}")
            return {new: Decimal(str(value[org])) for new, org in enumerate(self.original_index) if value[org] != 0}
        raise NonUniformSweep(f"spatial modulation {type(sm).__name__}")

    def _see(self, wf: ir.Waveform):
        self.duration = vmax(tofloat(wf.duration), self.duration)

    @staticmethod
    def _per_atom_sum(drives: "_Drives", targets: List[Dict[int, Any]], atom: int) -> ir.Waveform:
        """sum(target[atom] * wf ..., start=Constant(0.0, 0.0)) as the builder's `*` / `+` give it
        (waveform.py:90-135: every operator canonicalises its result)"""
        total: ir.Waveform = ir.Constant(0.0, 0.0)
        for (_, wf), tgt in zip(drives.items, targets):
            if atom in tgt:
                total = canonicalize(ir.Add(total, canonicalize(ir.Scale(tgt[atom], wf))))
        return total

    # -- emulator_ir.py:77-128 -----------------------------------------------------------------------------
    def construct_detuning(self, node) -> List[ir.DetuningTerm]:
        if node is None:
            return []
        drives = _Drives(node, self.env)
        terms = []
        if len(drives) <= self.n_atoms:
            for sm, wf in drives.items:
                self._see(wf)
                terms.append(ir.DetuningTerm(ir.DetuningOperatorData(self.target_atoms(sm)), wf))
            return terms
        targets = [self.target_atoms(sm) for sm, _ in drives.items]
        for atom in range(self.n_atoms):
            if not any(atom in t for t in targets):
                continue
            wf = self._per_atom_sum(drives, targets, atom)
            self._see(wf)
            terms.append(ir.DetuningTerm(ir.DetuningOperatorData({atom: Decimal("1.0")}), canonicalize(wf)))
        return terms

    # -- emulator_ir.py:130-262 ----------------------------------------------------------------------------
    def construct_rabi(self, amplitude, phase) -> List[ir.RabiTerm]:
        import bloqade.analog.ir.control.field as F

        terms: List[ir.RabiTerm] = []
        if amplitude is None:
            return terms
        amp = _Drives(amplitude, self.env)
        sym, asym = ir.RabiOperatorType.RabiSymmetric, ir.RabiOperatorType.RabiAsymmetric

        def shared_terms(op_type, rabi_phase):
            for sm, wf in amp.items:
                self._see(wf)
                tgt = self.target_atoms(sm)
                if len(tgt) == 0:
                    continue
                if len(tgt) == 1:
                    (scale,) = tgt.values()
                    if scale != 1:
                        wf = canonicalize(ir.Scale(scale, wf))
                terms.append(ir.RabiTerm(ir.RabiOperatorData(op_type, tgt), canonicalize(wf), rabi_phase))

        if phase is None and len(amp) <= self.n_atoms:
            shared_terms(sym, None)
            return terms
        if phase is None:
            targets = [self.target_atoms(sm) for sm, _ in amp.items]
            for atom in range(self.n_atoms):
                if not any(atom in t for t in targets):
                    continue
                wf = self._per_atom_sum(amp, targets, atom)
                self._see(wf)
                terms.append(ir.RabiTerm(ir.RabiOperatorData(sym, {atom: Decimal("1.0")}), canonicalize(wf), None))
            return terms
        ph = _Drives(phase, self.env)
        if len(phase.drives) == 1 and F.UniformModulation() in phase.drives and len(amp) <= self.n_atoms:
            ((_, phase_wf),) = ph.items
            self._see(phase_wf)
            shared_terms(asym, phase_wf)
            return terms
        ph_targets = [self.target_atoms(sm) for sm, _ in ph.items]
        amp_targets = [self.target_atoms(sm) for sm, _ in amp.items]
        for atom in range(self.n_atoms):
            if not any(atom in t for t in amp_targets):
                continue
            phase_wf = self._per_atom_sum(ph, ph_targets, atom)
            amp_wf = self._per_atom_sum(amp, amp_targets, atom)
            self._see(amp_wf)
            self._see(phase_wf)
            terms.append(ir.RabiTerm(ir.RabiOperatorData(asym, {atom: 1}), canonicalize(amp_wf), canonicalize(phase_wf)))
        return terms

    def visit_pulse(self, node) -> ir.Fields:
        import bloqade.analog.ir.control.pulse as P

        if type(node) is not P.Pulse:
            raise NonUniformSweep("pulse is not flat")
        detuning = node.fields.get(P.detuning)
        amplitude = node.fields.get(P.rabi.amplitude)
        phase = node.fields.get(P.rabi.phase)
        return ir.Fields(self.construct_detuning(detuning), self.construct_rabi(amplitude, phase))


class SweepProgram:
    """the whole sweep as ONE `ir.EmulatorProgram` over vector numbers + the per-task registers"""

    def __init__(self, circuit, assignments: Sequence[Dict[str, Any]], blockade_radius=0.0, use_hyperfine=False):
        if not assignments:
            raise NonUniformSweep("empty sweep")
        self.n_tasks = len(assignments)
        names = list(assignments[0].keys())
        env: Dict[str, Any] = {}
        for name in names:
            values = [a[name] for a in assignments]
            if isinstance(values[0], (list, tuple)):
                first = list(values[0])
                if any(list(v) != first for v in values[1:]):
                    raise NonUniformSweep("vector parameter varies over the sweep")
                env[name] = first
            else:
                env[name] = dec_vector(values)
        b = _Builder(circuit, env, blockade_radius, use_hyperfine)
        self.atom_type = ir.ThreeLevelAtom if b.is_hyperfine else ir.TwoLevelAtom
        self.blockade_radius = Decimal(str(blockade_radius))
        self.filling = [int(f) for f in b.filling]
        self._positions, self._sites = b.positions, b.sites
        self.duration = b.duration
        self.pulses = b.pulses
        self.n_atoms = b.n_atoms
        self.positions_vary = any(isinstance(c, DecVec) for pos in b.sites for c in pos)

    def register(self, k: int) -> ir.Register:
        """the emulator register of task k (emulator_ir.py:54-75)"""
        positions = [tuple(task_value(c, k) for c in pos) for pos in self._positions]
        sites = [tuple(float(task_value(c, k)) for c in pos) for pos in self._sites]
        return ir.Register(self.atom_type, positions, self.blockade_radius, ir.Geometry(sites, self.filling))

    def site_key(self, k: int):
        return tuple(tuple(task_value(c, k) for c in pos) for pos in self._sites)

    def sites_decimal(self, k: int):
        return [tuple(task_value(c, k) for c in pos) for pos in self._sites]

    def program(self, register: ir.Register) -> ir.EmulatorProgram:
        return ir.EmulatorProgram(register, self.duration, self.pulses)
