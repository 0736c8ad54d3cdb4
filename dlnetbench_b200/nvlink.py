"""NVLink byte counters of one GPU (NVML), read around a batch of launches.

north_star asks for NVLink GB/s per direction as evidence; `ncu` cannot wrap a multi-process
job, so the per-direction traffic of the multi-GPU kernels is taken from the driver's link
counters: NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_{TX,RX} (payload, KiB, summed over the 18 links)
and ..._RAW_{TX,RX} (payload + protocol overhead). delta / launches = bytes per launch.
"""
from __future__ import annotations

import time


class LinkCounters:
    def __init__(self, device_index: int):
        self.ok, self.why = False, ""
        try:
            import pynvml as nv

            nv.nvmlInit()
            self.nv = nv
            self.h = nv.nvmlDeviceGetHandleByIndex(device_index)
            self.fields = [("data_tx", nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX), ("data_rx", nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX),
                           ("raw_tx", nv.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_TX), ("raw_rx", nv.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_RX)]
            self.read()
            self.ok = True
        except Exception as e:  # no NVML, no NVLink, or counters not exposed on this box
            self.why = f"{type(e).__name__}: {e}"

    def read(self) -> dict:
        """Counters in bytes (scope UINT_MAX = sum over all links of the GPU)."""
        nv = self.nv
        vals = nv.nvmlDeviceGetFieldValues(self.h, [(fid, 0xFFFFFFFF) for _, fid in self.fields])
        out = {}
        for (name, _), v in zip(self.fields, vals):
            if v.nvmlReturn != nv.NVML_SUCCESS:
                raise RuntimeError(f"NVML field {name}: return code {v.nvmlReturn}")
            t = v.valueType
            val = (v.value.ullVal if t == nv.NVML_VALUE_TYPE_UNSIGNED_LONG_LONG else v.value.ulVal if t == nv.NVML_VALUE_TYPE_UNSIGNED_LONG
                   else v.value.uiVal if t == nv.NVML_VALUE_TYPE_UNSIGNED_INT else int(v.value.dVal))
            out[name] = int(val) * 1024
        return out

    def settle(self) -> dict:
        """Read after the counters had time to catch up with finished work."""
        time.sleep(0.25)
        return self.read()


def per_launch(before: dict, after: dict, launches: int) -> dict:
    return {k + "_bytes_per_launch": (after[k] - before[k]) / max(1, launches) for k in before}
