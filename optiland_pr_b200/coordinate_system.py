"""CoordinateSystem: decenter + Euler tilt holder (optiland/coordinate_system.py:28-71).

`localize` / `globalize` (coordinate_system.py:73-107) run inside the kernel
(csrc/trace_common.cuh: frame_localize / frame_globalize); this class only carries the six
numbers and the optional `reference_cs` chain. Values may be Python floats or torch tensors
(tensors with requires_grad make the parameter differentiable).
"""
from __future__ import annotations


class CoordinateSystem:
    def __init__(self, x=0, y=0, z=0, rx=0, ry=0, rz=0, reference_cs=None):
        self.x = x
        self.y = y
        self.z = z
        self.rx = rx
        self.ry = ry
        self.rz = rz
        self.reference_cs = reference_cs

    def to_dict(self) -> dict:
        return {"x": float(self.x), "y": float(self.y), "z": float(self.z), "rx": float(self.rx),
                "ry": float(self.ry), "rz": float(self.rz),
                "reference_cs": self.reference_cs.to_dict() if self.reference_cs else None}

    @classmethod
    def from_dict(cls, data: dict) -> "CoordinateSystem":
        ref = cls.from_dict(data["reference_cs"]) if data.get("reference_cs") else None
        return cls(data.get("x", 0), data.get("y", 0), data.get("z", 0), data.get("rx", 0),
                   data.get("ry", 0), data.get("rz", 0), ref)
