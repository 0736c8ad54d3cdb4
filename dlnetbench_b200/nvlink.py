"""NVLink byte counters of one GPU (NVML), read around a batch of launches.

north_star asks for NVLink GB/s per direction as evidence; `ncu` cannot wrap a multi-process
job, so the per-direction traffic of the multi-GPU kernels is taken from the driver's link
counters: NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_{TX,RX} (payload, KiB, summed over the 18 links)
and ..._RAW_{TX,RX} (payload + protocol overhead). delta / launches = bytes per launch.
"""
from __future__ import annotations

import time


class LinkCounters:
    """Tries, in order: NVML field values summed by the driver (scope UINT_MAX), the same fields per link (scope =
    link 0..17) summed here, and `nvidia-smi nvlink -gt d` (the tool's own reading of the data counters)."""

    def __init__(self, device_index: int):
        self.ok, self.why, self.how, self.index = False, "", "", device_index
        errs = []
        for how in ("nvml_sum", "nvml_per_link", "nvidia_smi"):
            try:
                self.how = how
                if how != "nvidia_smi":
                    import pynvml as nv

                    nv.nvmlInit()
                    self.nv = nv
                    self.h = nv.nvmlDeviceGetHandleByIndex(device_index)
                    self.fields = [("data_tx", nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX), ("data_rx", nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX),
                                   ("raw_tx", nv.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_TX), ("raw_rx", nv.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_RX)]
                first = self.read()
                if not any(first.values()) and how != "nvidia_smi":
                    raise RuntimeError("all counters read zero")
                self.ok = True
                return
            except Exception as e:  # no NVML, no NVLink, or counters not exposed this way on this box
                errs.append(f"{how}: {type(e).__name__}: {e}")
        self.why = "; ".join(errs)

    def _fields(self, scope) -> dict:
        nv = self.nv
        vals = nv.nvmlDeviceGetFieldValues(self.h, [(fid, scope) for _, fid in self.fields])
        out = {}
        for (name, _), v in zip(self.fields, vals):
            if v.nvmlReturn != nv.NVML_SUCCESS:
                raise RuntimeError(f"NVML field {name} scope {scope}: return code {v.nvmlReturn}")
            t = v.valueType
            val = (v.value.ullVal if t == nv.NVML_VALUE_TYPE_UNSIGNED_LONG_LONG else v.value.ulVal if t == nv.NVML_VALUE_TYPE_UNSIGNED_LONG
                   else v.value.uiVal if t == nv.NVML_VALUE_TYPE_UNSIGNED_INT else int(v.value.dVal))
            out[name] = int(val) * 1024
        return out

    def read(self) -> dict:
        """Counters in bytes, summed over the links of the GPU."""
        if self.how == "nvml_sum":
            return self._fields(0xFFFFFFFF)
        if self.how == "nvml_per_link":
            tot, seen = {}, 0
            for link in range(18):
                try:
                    one = self._fields(link)
                except RuntimeError:
                    continue
                seen += 1
                for k, v in one.items():
                    tot[k] = tot.get(k, 0) + v
            if not seen:
                raise RuntimeError("no link answered")
            return tot
        import re
        import subprocess

        txt = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(self.index)], capture_output=True, text=True, timeout=20).stdout
        txs, rxs = re.findall(r"Data Tx:\s*(\d+)\s*KiB", txt), re.findall(r"Data Rx:\s*(\d+)\s*KiB", txt)
        if not txs or not rxs:  # this pool prints "Data Tx: N/A" for every link: the counters are not exposed
            raise RuntimeError("nvidia-smi nvlink -gt d shows no readable counter: " + " ".join(txt.split())[:100])
        return {"data_tx": sum(int(x) for x in txs) * 1024, "data_rx": sum(int(x) for x in rxs) * 1024}

    def settle(self) -> dict:
        """Read after the counters had time to catch up with finished work."""
        time.sleep(0.25)
        return self.read()


def per_launch(before: dict, after: dict, launches: int) -> dict:
    return {k + "_bytes_per_launch": (after[k] - before[k]) / max(1, launches) for k in before}
