"""Host-side mirror of the reference's ``internal/hal`` Driver/Device interfaces
(internal/hal/device.go:155-198) for a CUDA driver that, unlike GPULinuxDriver
(internal/hal/gpu_linux.go:20-31), reports ``SHA256d: true`` because it comes
with a compute path.  Enumeration uses the CUDA runtime, not /sys/class/drm.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import List

from .capi import DeviceInfo, check, lib

FAMILY_GPU = "gpu"


@dataclass(frozen=True)
class Identity:
    """hal.Identity, device.go:88-132."""
    id: str
    family: str
    vendor: str
    model: str

    def validate(self) -> None:
        if not self.id:
            raise ValueError("hal: Identity.ID must not be empty")
        if self.family not in ("asic", "gpu", "cpu"):
            raise ValueError(f"hal: Identity.Family {self.family!r} is not a valid Family")
        for ch in self.id:
            if ch in " \t\n/":
                raise ValueError(f"hal: Identity.ID contains forbidden character {ch!r}")


@dataclass(frozen=True)
class Capabilities:
    """hal.Capabilities, device.go:140-148."""
    sha256d: bool
    general_compute: bool


class CudaDevice:
    """hal.Device for one CUDA GPU."""

    def __init__(self, index: int, info: DeviceInfo):
        self.index = index
        self.info = info
        self._shutdown = False

    def identity(self) -> Identity:
        return Identity(self.info.id.decode(), FAMILY_GPU, self.info.vendor.decode(), self.info.model.decode())

    def capabilities(self) -> Capabilities:
        return Capabilities(bool(self.info.cap_sha256d), bool(self.info.cap_general_compute))

    def telemetry(self):
        """(power_w, sm_clock_mhz, temperature_c) via NVML — feeds joules-per-terahash."""
        from .capi import CTelemetry
        t = CTelemetry()
        check(lib().otd_device_telemetry(self.index, ctypes.byref(t)))
        return t.power_w, t.sm_clock_mhz, t.temperature_c

    def shutdown(self) -> None:
        """Idempotent, like the reference requires (device.go:165-169)."""
        self._shutdown = True


class CudaDriver:
    """hal.Driver: Name() == "cuda"; Enumerate lists the visible CUDA devices."""

    def name(self) -> str:
        return "cuda"

    def enumerate(self) -> List[CudaDevice]:
        n = check(lib().otd_device_count())
        devs = []
        for i in range(n):
            info = DeviceInfo()
            check(lib().otd_device_info(i, ctypes.byref(info)))
            devs.append(CudaDevice(i, info))
        return devs
