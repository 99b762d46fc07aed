import os, sys, time, torch, torch.distributed as dist
rank = int(os.environ['RANK']); world = int(os.environ['WORLD_SIZE']); local = int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local); dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
ok = {}
try:
    import torch.distributed._symmetric_memory as symm
    buf = symm.empty(1 << 20, dtype=torch.uint8, device=dev)
    hdl = symm.rendezvous(buf, dist.group.WORLD.group_name)
    peer0 = hdl.get_buffer(0, (1 << 20,), torch.uint8)
    peer0[rank * 16:(rank + 1) * 16] = rank + 1
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    if rank == 0:
        print('symm ok', buf[:16 * world:16].tolist(), flush=True)
    ok['symm'] = True
except Exception as e:
    print(rank, 'symm failed', repr(e)[:300], flush=True)
try:
    t = torch.zeros(1 << 20, dtype=torch.uint8, device=dev)
    st = t.untyped_storage()
    h = st._share_cuda_()
    objs = [None] * world
    dist.all_gather_object(objs, h if rank == 0 else None)
    if rank != 0:
        h0 = objs[0]
        st0 = torch.UntypedStorage._new_shared_cuda(*h0)
        p0 = torch.empty(0, dtype=torch.uint8, device=torch.device('cuda', h0[0])).set_(st0)
        print(rank, 'ipc tensor device', p0.device, flush=True)
        src = torch.full((16,), rank + 1, dtype=torch.uint8, device=dev)
        p0[rank * 16:(rank + 1) * 16].copy_(src)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    if rank == 0:
        print('ipc ok', t[:16 * world:16].tolist(), flush=True)
except Exception as e:
    print(rank, 'ipc failed', repr(e)[:300], flush=True)
dist.destroy_process_group()
