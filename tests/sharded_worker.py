"""One rank of the multi-GPU parity test (launched by tests/test_multi_gpu_hw.py under torchrun):
a global batch goes through rtldavis_b200.shard.ShardedDemodulator on WORLD_SIZE GPUs, rank 0
compares the gathered records with the C oracle on the whole batch and prints OK / MISMATCH."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np
import torch
import torch.distributed as dist

from oracle import c_oracle as CO
from rtldavis_b200 import protocol, shard, synth


def rows(rec):
    return [(int(r["stream"]), int(r["call"]), int(r["index"]), r["data"].tobytes(), bool(r["crc_ok"]))
            for r in rec]


def main():
    rank, world, local = shard.rank_world()
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n_streams, n_blocks = 37, 24                     # 37: uneven over 2, 4 and 8 ranks
    iq, _ = synth.make_batch(n_streams, n_blocks * 8192, 3, 15.0, seed0=3100, df_range_hz=2500.0, ppm_range=40.0)
    ok = True
    for weights in (None, [1.0 + 0.5 * (r % 2) for r in range(world)]):
        sd = shard.ShardedDemodulator(protocol.new_packet_config(14), n_streams, max_blocks=10, weights=weights)
        got = []
        for k0, k1 in ((0, 10), (10, 17), (17, 24)):         # carried state across calls
            part = iq[sd.first_stream: sd.first_stream + sd.count, 2 * 8192 * k0: 2 * 8192 * k1]
            sd.submit(np.ascontiguousarray(part))
            g = sd.fetch_gathered()
            if rank == 0:
                got += rows(g)
        sd.close()
        if rank == 0:
            want = rows(CO.batch_records(iq))
            ok = ok and sorted(got) == sorted(want) and len(want) > 100
    if rank == 0:
        print("SHARDED", "OK" if ok else "MISMATCH", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
