"""Bridge from a real `bloqade.analog` installation to the B200 backend IR.

Only imported when `bloqade.analog` is importable (a user's environment, or the build container
through baseline/refshim.py).  Converts the reference's `EmulatorProgram`
(`src/bloqade/analog/emulate/ir/emulator.py:241-246`) - register, per-coupling `Fields`, and the
`JITWaveform`s - into `bloqade_analog_b200.ir`, node for node.  Waveforms are taken from
`JITWaveform.canonicalized_ir` (emulator.py:32-63), i.e. after assignment and after `Sample`
nodes have been rewritten into Append-of-Linear/Constant, which is exactly what the reference's
default "interpret" runtime evaluates (emulator.py:71-72).
"""
from __future__ import annotations

from decimal import Decimal

from . import ir


def convert_waveform(node) -> ir.Waveform:
    import bloqade.analog.ir.control.waveform as W

    if isinstance(node, W.Linear):
        return ir.Linear(node.start(), node.stop(), node.duration())
    if isinstance(node, W.Constant):
        return ir.Constant(node.value(), node.duration())
    if isinstance(node, W.Poly):
        return ir.Poly([c() for c in node.coeffs], node.duration())
    if isinstance(node, W.PythonFn):
        kwargs = {p.name: float(p(**node.default_param_values)) for p in node.parameters}
        fn = node.fn
        return ir.PythonFn(lambda t, _fn=fn, _kw=kwargs: _fn(t, **_kw), node.duration(**node.default_param_values))
    if isinstance(node, W.Smooth):
        return ir.PythonFn(lambda t, _n=node: float(_n.eval_decimal(Decimal(str(t)))), node.duration())
    if isinstance(node, W.Append):
        return ir.Append([convert_waveform(w) for w in node.waveforms])
    if isinstance(node, W.Slice):
        inner = convert_waveform(node.waveform)
        return ir.Slice(inner, node.start(), node.stop())
    if isinstance(node, W.Negative):
        return ir.Negative(convert_waveform(node.waveform))
    if isinstance(node, W.Scale):
        return ir.Scale(node.scalar(), convert_waveform(node.waveform))
    if isinstance(node, W.Add):
        return ir.Add(convert_waveform(node.left), convert_waveform(node.right))
    if isinstance(node, W.Record):
        return convert_waveform(node.waveform)
    if isinstance(node, W.Sample):
        from bloqade.analog.compiler.rewrite.python.waveform import NormalizeWaveformPython

        return convert_waveform(NormalizeWaveformPython().visit(node))
    raise TypeError(f"unsupported waveform node {type(node).__name__}")


def convert_jit_waveform(jit) -> ir.Waveform:
    return convert_waveform(jit.canonicalized_ir)


def convert_register(reg) -> ir.Register:
    from bloqade.analog.emulate.ir.atom_type import ThreeLevelAtomType

    atom_type = ir.ThreeLevelAtom if isinstance(reg.atom_type, ThreeLevelAtomType) else ir.TwoLevelAtom
    geometry = None
    if reg.geometry is not None:
        geometry = ir.Geometry(
            [(float(x), float(y)) for x, y in reg.geometry.sites], [int(f) for f in reg.geometry.filling]
        )
    return ir.Register(atom_type, list(reg.sites), reg.blockade_radius, geometry)


def convert_program(program) -> ir.EmulatorProgram:
    """reference EmulatorProgram -> bloqade_analog_b200.ir.EmulatorProgram"""
    pulses = {}
    cache = {}

    def wf(jit):
        if jit is None:
            return None
        key = id(jit)
        if key not in cache:
            cache[key] = convert_jit_waveform(jit)
        return cache[key]

    for coupling, fields in program.pulses.items():
        det = [
            ir.DetuningTerm(ir.DetuningOperatorData(dict(t.operator_data.target_atoms)), wf(t.amplitude))
            for t in fields.detuning
        ]
        rabi = [
            ir.RabiTerm(
                ir.RabiOperatorData(
                    ir.RabiOperatorType(int(t.operator_data.operator_type)), dict(t.operator_data.target_atoms)
                ),
                wf(t.amplitude),
                wf(t.phase),
            )
            for t in fields.rabi
        ]
        pulses[ir.LevelCoupling(coupling.value)] = ir.Fields(det, rabi)
    return ir.EmulatorProgram(convert_register(program.register), float(program.duration), pulses)
