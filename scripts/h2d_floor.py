#!/usr/bin/env python3
"""Host -> device floor under the e2e legs of bench.py: every rank copies, concurrently with all the others,
what one e2e step of the default workload moves from pinned host memory:

  strided   the confidence planes only (one cudaMemcpy2DAsync per tensor, exactly the library's call: 1 of 5
            cif planes, 1 of 9 caf planes), 60 MB per 64 images -- what host_transfer='mapped' copies by DMA
            (strided_torch: the same bytes through torch's strided copy_, for comparison);
  plain     the same byte count as one contiguous copy;
  full      every plane (430 MB per 64 images) -- what host_transfer='copy' moves.

Run under torchrun at N = 1, 2, 4, 8; rank 0 prints one JSON line with the aggregate GB/s (bytes of all
ranks / max-over-ranks time).  No decode runs here: this is the transfer alone."""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)
        import bench
        bench.bind_to_gpu_numa(local)
    B, H = 64, 81
    h_cif = torch.rand((B, 17, 5, H, H)).pin_memory()
    h_caf = torch.rand((B, 19, 9, H, H)).pin_memory()
    d_cif, d_caf = torch.empty_like(h_cif, device=dev), torch.empty_like(h_caf, device=dev)
    n_conf = (h_cif[:, :, 0].numel() + h_caf[:, :, 0].numel()) * 4
    h_plain = torch.rand((n_conf // 4,)).pin_memory()
    d_plain = torch.empty_like(h_plain, device=dev)

    # the library's copy of the confidence planes: one cudaMemcpy2DAsync per tensor (csrc/api.cu,
    # copy_conf_planes): rows = (image, field) blocks, pitch = all planes of a block, width = one plane
    from cuda.bindings import runtime as cudart
    stream = torch.cuda.current_stream().cuda_stream
    cells = H * H * 4

    def strided():
        for d, h, planes, blocks in ((d_cif, h_cif, 5, B * 17), (d_caf, h_caf, 9, B * 19)):
            err, = cudart.cudaMemcpy2DAsync(d.data_ptr(), planes * cells, h.data_ptr(), planes * cells, cells, blocks,
                                            cudart.cudaMemcpyKind.cudaMemcpyHostToDevice, stream)
            assert int(err) == 0, err

    def strided_torch():
        d_cif[:, :, 0].copy_(h_cif[:, :, 0], non_blocking=True)
        d_caf[:, :, 0].copy_(h_caf[:, :, 0], non_blocking=True)

    def plain():
        d_plain.copy_(h_plain, non_blocking=True)

    def full():
        d_cif.copy_(h_cif, non_blocking=True)
        d_caf.copy_(h_caf, non_blocking=True)

    out = {}
    for name, fn, nbytes, reps in (('strided', strided, n_conf, 50), ('strided_torch', strided_torch, n_conf, 20),
                                   ('plain', plain, n_conf, 50),
                                   ('full', full, h_cif.numel() * 4 + h_caf.numel() * 4, 10)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        dt = float(dt.item())
        out[name] = {'bytes_per_rank_per_step': nbytes, 'ms_per_step': 1e3 * dt / reps,
                     'aggregate_gbs': world * nbytes * reps / dt / 1e9,
                     'images_per_s_floor': world * B * reps / dt}
    if rank == 0:
        print(json.dumps({'what': 'concurrent pinned host -> device copies of one e2e step per rank', 'n_gpus': world,
                          'host_cpus': os.cpu_count(), **out}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
