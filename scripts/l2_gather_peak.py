"""Measures the random-gather roof of this GPU for the hash-grid kernels and writes
profiles/l2_gather_peak.json (via gpurun_out/): 4-, 8- and 16-byte requests at uniformly random entries of a 24.4 MB table
(the fp16 hash table: L2-resident), of a 2 MB table (one hashed level) and of a 1 GB table (HBM-resident, for contrast).
Requests/s and bytes/s = requested bytes (not sectors).  CUDA events, best of 10 after 3 warm-ups."""
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arcnerf_b200 import _lib   # noqa: E402


def main():
    lib = _lib.load()
    lib.arcb200_probe_gather.restype = C.c_longlong
    n_threads, trips = 148 * 2048, 64
    sink = torch.empty(n_threads, dtype=torch.int32, device='cuda')
    out = {'device': torch.cuda.get_device_name(0), 'threads': n_threads, 'loads_in_flight_per_thread': 8, 'cases': []}
    for label, nbytes in (('2 MB (one hashed level, fp16)', 2 << 20), ('24.4 MB (whole fp16 hash table)', 24392480),
                          ('1 GB (HBM resident)', 1 << 30)):
        table = torch.empty(nbytes, dtype=torch.uint8, device='cuda').random_(0, 255)
        for req in (4, 8, 16):
            best = 1e9
            for it in range(13):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                n = lib.arcb200_probe_gather(C.c_void_p(table.data_ptr()), C.c_int64(nbytes), req, n_threads, trips, it * 7919,
                                             C.c_void_p(sink.data_ptr()), C.c_void_p(_lib.current_stream()))
                b.record()
                torch.cuda.synchronize()
                assert n > 0, _lib.last_error()
                if it >= 3:
                    best = min(best, a.elapsed_time(b))
            out['cases'].append(dict(table=label, table_bytes=nbytes, request_bytes=req, requests=int(n), ms=best,
                                     giga_requests_per_s=n / best / 1e6, requested_GBps=n * req / best / 1e6,
                                     sector_GBps=n * 32 / best / 1e6))
        del table
    os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
    with open(os.path.join(ROOT, 'gpurun_out', 'l2_gather_peak.json'), 'w') as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == '__main__':
    main()
