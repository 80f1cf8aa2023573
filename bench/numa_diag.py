import os, glob, torch, time
print('affinity', len(os.sched_getaffinity(0)), sorted(os.sched_getaffinity(0))[:8], '...')
for n in sorted(glob.glob('/sys/devices/system/node/node*')):
    print(n, open(n + '/cpulist').read().strip())
p = torch.cuda.get_device_properties(0)
bus = '%04x:%02x:%02x.0' % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
try: print('gpu', bus, 'numa', open('/sys/bus/pci/devices/%s/numa_node' % bus).read().strip(), 'local_cpulist', open('/sys/bus/pci/devices/%s/local_cpulist' % bus).read().strip())
except Exception as e: print('no sysfs', e)
def bw(tag):
    h = torch.empty(64 << 20, dtype=torch.uint8).pin_memory(); d = torch.empty_like(h, device='cuda')
    for _ in range(3): d.copy_(h, non_blocking=True)
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): d.copy_(h, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print(tag, 'H2D GB/s', 20 * h.numel() / e0.elapsed_time(e1) / 1e6)
bw('default')
for n in sorted(glob.glob('/sys/devices/system/node/node*')):
    cl = open(n + '/cpulist').read().strip(); cpus = set()
    for part in cl.split(','):
        a, _, b = part.partition('-'); cpus |= set(range(int(a), int(b or a) + 1))
    try:
        os.sched_setaffinity(0, cpus & os.sched_getaffinity(0) or cpus); bw('bound to ' + os.path.basename(n))
    except Exception as e: print('bind failed', n, e)
