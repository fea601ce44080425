"""JSON (de)serialisation of the backend IR - keeps tasks picklable/serialisable like the reference's
`Serializer` registry (src/bloqade/analog/serialize.py, task/bloqade.py:91-106) and stores the golden
programs under tests/golden/.  Decimals travel as strings; `PythonFn` is stored as its flattened
piece table (a host callable cannot be serialised)."""
from __future__ import annotations

from decimal import Decimal
from typing import Any, Dict

from . import ir


class PieceTable(ir.Waveform):
    """A waveform given directly by its flattened pieces (deserialised PythonFn / Smooth)."""

    def __init__(self, pieces, duration):
        self.pieces = [ir.Piece(*p[:4], tuple(p[4])) for p in pieces]
        self.duration = ir._dec(duration)

    def eval_decimal(self, clock_s):
        return Decimal(str(ir.eval_pieces(self.pieces, float(clock_s))))

    def __call__(self, clock_s):
        return ir.eval_pieces(self.pieces, float(clock_s))


def waveform_to_dict(w: ir.Waveform) -> Dict[str, Any]:
    s = str
    if isinstance(w, ir.Linear):
        return {"k": "linear", "start": s(w.start), "stop": s(w.stop), "duration": s(w.duration)}
    if isinstance(w, ir.Constant):
        return {"k": "constant", "value": s(w.value), "duration": s(w.duration)}
    if isinstance(w, ir.Poly):
        return {"k": "poly", "coeffs": [s(c) for c in w.coeffs], "duration": s(w.duration)}
    if isinstance(w, ir.Append):
        return {"k": "append", "waveforms": [waveform_to_dict(x) for x in w.waveforms]}
    if isinstance(w, ir.Slice):
        return {"k": "slice", "waveform": waveform_to_dict(w.waveform), "start": s(w.start), "stop": s(w.stop)}
    if isinstance(w, ir.Negative):
        return {"k": "negative", "waveform": waveform_to_dict(w.waveform)}
    if isinstance(w, ir.Scale):
        return {"k": "scale", "scalar": s(w.scalar), "waveform": waveform_to_dict(w.waveform)}
    if isinstance(w, ir.Add):
        return {"k": "add", "left": waveform_to_dict(w.left), "right": waveform_to_dict(w.right)}
    if isinstance(w, (ir.PythonFn, PieceTable)):
        pieces = w.pieces if isinstance(w, PieceTable) else ir.flatten_waveform(w)
        return {"k": "pieces", "duration": s(w.duration),
                "pieces": [[p.lo if p.lo != float("-inf") else "-inf", p.hi, p.offset, p.scale, list(p.coeffs)] for p in pieces]}
    raise TypeError(type(w).__name__)


def waveform_from_dict(d: Dict[str, Any]) -> ir.Waveform:
    k = d["k"]
    D = Decimal
    if k == "linear":
        return ir.Linear(D(d["start"]), D(d["stop"]), D(d["duration"]))
    if k == "constant":
        return ir.Constant(D(d["value"]), D(d["duration"]))
    if k == "poly":
        return ir.Poly([D(c) for c in d["coeffs"]], D(d["duration"]))
    if k == "append":
        return ir.Append([waveform_from_dict(x) for x in d["waveforms"]])
    if k == "slice":
        return ir.Slice(waveform_from_dict(d["waveform"]), D(d["start"]), D(d["stop"]))
    if k == "negative":
        return ir.Negative(waveform_from_dict(d["waveform"]))
    if k == "scale":
        return ir.Scale(D(d["scalar"]), waveform_from_dict(d["waveform"]))
    if k == "add":
        return ir.Add(waveform_from_dict(d["left"]), waveform_from_dict(d["right"]))
    if k == "pieces":
        pieces = [[float(p[0]), p[1], p[2], p[3], p[4]] for p in d["pieces"]]
        return PieceTable(pieces, D(d["duration"]))
    raise ValueError(k)


def program_to_dict(p: ir.EmulatorProgram) -> Dict[str, Any]:
    reg = p.register
    pulses = {}
    for coupling, fields in p.pulses.items():
        pulses[coupling.value] = {
            "detuning": [
                {"target_atoms": {str(a): str(v) for a, v in t.operator_data.target_atoms.items()},
                 "amplitude": waveform_to_dict(t.amplitude)}
                for t in fields.detuning
            ],
            "rabi": [
                {"operator_type": int(t.operator_data.operator_type),
                 "target_atoms": {str(a): str(v) for a, v in t.operator_data.target_atoms.items()},
                 "amplitude": waveform_to_dict(t.amplitude),
                 "phase": None if t.phase is None else waveform_to_dict(t.phase)}
                for t in fields.rabi
            ],
        }
    return {
        "register": {
            "n_level": reg.atom_type.n_level,
            "sites": [[str(x), str(y)] for x, y in reg.sites],
            "blockade_radius": str(reg.blockade_radius),
            "geometry": {"sites": [list(s) for s in reg.geometry.sites], "filling": list(reg.geometry.filling)},
        },
        "duration": p.duration,
        "pulses": pulses,
    }


def program_from_dict(d: Dict[str, Any]) -> ir.EmulatorProgram:
    r = d["register"]
    atom_type = ir.ThreeLevelAtom if r["n_level"] == 3 else ir.TwoLevelAtom
    geometry = ir.Geometry([tuple(s) for s in r["geometry"]["sites"]], list(r["geometry"]["filling"]))
    reg = ir.Register(atom_type, [(Decimal(x), Decimal(y)) for x, y in r["sites"]], Decimal(r["blockade_radius"]), geometry)
    pulses = {}
    for cname, f in d["pulses"].items():
        det = [
            ir.DetuningTerm(ir.DetuningOperatorData({int(a): Decimal(v) for a, v in t["target_atoms"].items()}),
                            waveform_from_dict(t["amplitude"]))
            for t in f["detuning"]
        ]
        rabi = [
            ir.RabiTerm(
                ir.RabiOperatorData(ir.RabiOperatorType(t["operator_type"]),
                                    {int(a): Decimal(v) for a, v in t["target_atoms"].items()}),
                waveform_from_dict(t["amplitude"]),
                None if t["phase"] is None else waveform_from_dict(t["phase"]),
            )
            for t in f["rabi"]
        ]
        pulses[ir.LevelCoupling(cname)] = ir.Fields(det, rabi)
    return ir.EmulatorProgram(reg, float(d["duration"]), pulses)
