"""what does NVML report for the NVLink throughput counters on this box? (diagnostic)"""
import pynvml as nv
nv.nvmlInit()
h = nv.nvmlDeviceGetHandleByIndex(0)
ids = {"DATA_TX": nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, "DATA_RX": nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX,
       "RAW_TX": nv.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_TX, "RAW_RX": nv.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_RX}
for scope in (None, 0, 0xFFFFFFFF):
    for name, fid in ids.items():
        try:
            arg = [fid] if scope is None else [(fid, scope)]
            v = nv.nvmlDeviceGetFieldValues(h, arg)[0]
            print("scope", scope, name, "ret", v.nvmlReturn, "type", v.valueType, "ull", v.value.ullVal)
        except Exception as e:
            print("scope", scope, name, "exception", repr(e))
for link in range(2):
    try:
        print("link", link, "state", nv.nvmlDeviceGetNvLinkState(h, link))
    except Exception as e:
        print("link", link, "state exception", repr(e))
    for kind in (0, 1):
        try:
            print("  utilization counter", kind, nv.nvmlDeviceGetNvLinkUtilizationCounter(h, link, kind))
        except Exception as e:
            print("  utilization counter", kind, "exception", repr(e))
