"""Experiment helper: wall time of msbfs_load_graph_file (read -> pinned staging -> H2D -> CSR
build on the device) on a graph.bin in /dev/shm, next to msbfs_set_graph_edges (host array)
and msbfs_set_graph_edges_device.  usage: python tools/loader_bench.py [scale]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft
import torch

pkg = graft.load_package()
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 24
n, m = 1 << scale, 16 << scale
buf = torch.empty((m, 2), dtype=torch.int32, device="cuda:0")
assert pkg.lib().msbfs_gen_edges_device(0, pkg.GEN_KRON, scale, 0, m, 1, buf.data_ptr()) == 0
uv = buf.cpu().numpy()
path = "/dev/shm/msbfs_loader_bench.bin"
pkg.write_graph_file(path, n, uv)
size_gb = os.path.getsize(path) / 1e9
with pkg.Engine(0) as eng:
    def with_threads(t):
        eng.set_option("stage_threads", t)
        eng.load_graph_file(path)
    for label, fn in [("load_graph_file (/dev/shm)", lambda: eng.load_graph_file(path)),
                      ("load_graph_file again", lambda: eng.load_graph_file(path)),
                      ("  stage_threads=2", lambda: with_threads(2)), ("  stage_threads=4", lambda: with_threads(4)),
                      ("  stage_threads=8", lambda: with_threads(8)), ("  stage_threads=12", lambda: with_threads(12)),
                      ("  stage_threads=16", lambda: with_threads(16)), ("  stage_threads=0 (default)", lambda: with_threads(0)),
                      ("set_graph_edges (host array)", lambda: eng.set_graph_edges(n, uv)),
                      ("set_graph_edges_device", lambda: eng.set_graph_edges_device(n, m, buf.data_ptr()))]:
        t0 = time.time(); fn(); dt = time.time() - t0
        print(f"{label:32s} {dt:7.3f} s  ({size_gb / dt:6.2f} GB/s of edge records; device build {eng.stats()['build_ms']:.0f} ms)", flush=True)
os.remove(path)
