"""Random-row read rate of HBM3e vs row size (tools/microbench/random_rows.cu): the ceiling of the edge walks,
which gather one 256-byte row per table per edge from tables far larger than L2.
usage (GPU box): python tools/bench_random_rows.py"""
import ctypes, os, sys
import torch

here = os.path.dirname(os.path.abspath(__file__))
lib = ctypes.CDLL(os.path.join(here, "microbench", "librandom_rows.so"))
lib.random_rows.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                            ctypes.c_int, ctypes.c_void_p]
dev = torch.device("cuda")
table_bytes = 2 << 30
table = torch.empty(table_bytes // 4, device=dev, dtype=torch.float32).uniform_()
sink = torch.zeros(4, device=dev)
num_pos = 42_200_000
flush = torch.zeros(64 << 20, device=dev)
stream = torch.cuda.current_stream().cuda_stream
print(f"table {table_bytes / 1e9:.2f} GB, {num_pos} random rows per launch, L2 flushed between launches")
for row_bytes in (64, 128, 256, 512):
    rows = table_bytes // row_bytes
    idx = torch.randint(0, rows, (num_pos,), device=dev, dtype=torch.int32)
    for unroll in (1, 2, 4, 8):
        for ctas_per_sm in (4, 8):
            blocks = 148 * ctas_per_sm
            best = 1e9
            for rep in range(4):
                flush.add_(1.0)
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                rc = lib.random_rows(table.data_ptr(), idx.data_ptr(), num_pos, row_bytes, sink.data_ptr(), unroll, blocks, stream)
                b.record()
                torch.cuda.synchronize()
                assert rc == 0, rc
                if rep:
                    best = min(best, a.elapsed_time(b))
            print(f"row {row_bytes:4d} B  unroll {unroll}  {ctas_per_sm} CTAs/SM: {best:7.3f} ms  {num_pos / best / 1e6:7.2f} G rows/s  "
                  f"{num_pos * row_bytes / best / 1e6:7.1f} GB/s (+{num_pos * 4 / best / 1e6:.0f} GB/s of indices)", flush=True)
