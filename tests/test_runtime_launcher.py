import os

import pytest
import torch
import torch.distributed as dist

from pytorch_distributed_b200 import launcher, runtime
from pytorch_distributed_b200.config import multi_node_addressing
from pytorch_distributed_b200.models import resnet50
from pytorch_distributed_b200.utils import ParamArena

from . import _mp_helpers as H


def test_bucket_plan_matches_torch_assignment():
    m = resnet50()
    params = list(m.parameters())
    # torch's planner on the reversed parameter list with the 1 MiB / 25 MiB policy
    idx, _ = dist._compute_bucket_assignment_by_size(params[::-1], [1 << 20, 25 << 20], [False] * len(params))
    want = [sorted(len(params) - 1 - i for i in b) for b in idx]
    a = ParamArena(m)
    ranges, pb = runtime.plan_buckets([e.offset for e in a.entries], [e.numel for e in a.entries])
    got = [sorted(i for i, b in enumerate(pb) if b == k) for k in range(len(ranges))]
    assert got == want
    assert ranges == a.bucket_ranges()


def test_reducer_releases_in_order():
    r = runtime.Reducer([2, 2, 1, 1, 0, 0], 3)
    assert r.mark_ready(0) == [] and r.mark_ready(1) == []      # bucket 2 complete but must wait
    assert r.mark_ready(5) == [] and r.mark_ready(4) == [0]
    assert r.mark_ready(2) == [] and r.mark_ready(3) == [1, 2] and r.done()
    with pytest.raises(RuntimeError):
        r.mark_ready(3)
    r.reset()
    assert not r.done() and r.pending(0) == 2
    r.mark_ready(4)
    assert r.flush() == [0, 1, 2] and r.done()


def test_numa_binding_is_safe():
    before = runtime.affinity_count()
    assert before >= 1
    assert runtime.bind_numa_for_pci("ffff:ff:1f.7") == -1       # unknown device: nothing changes
    assert runtime.affinity_count() == before
    node = runtime.bind_numa_node(0)
    assert node in (0, -1) and runtime.affinity_count() >= 1


def test_multi_node_addressing():
    env = {"MASTER_IP": "10.0.0.1", "MASTER_PORT": "1234", "WORLD_SIZE": "2", "RANK": "1"}
    assert multi_node_addressing(3, 4, env) == ("tcp://10.0.0.1:1234", 8, 7, 1)      # ref:restnet_ddp.py:87-94
    assert multi_node_addressing(0, 8, {}) == ("tcp://127.0.0.1:29500", 8, 0, 0)


def test_spawn_runs_all(tmp_path):
    launcher.spawn(H.child_ok, args=(str(tmp_path),), nprocs=3, bind_numa=True)
    assert sorted(os.listdir(tmp_path)) == ["ok0", "ok1", "ok2"]


def test_spawn_propagates_exception():
    with pytest.raises(launcher.ProcessRaisedException) as ei:
        launcher.spawn(H.child_raises, nprocs=2, grace_period=5)
    assert "boom from rank 1" in str(ei.value) and ei.value.error_index == 1


def test_spawn_reports_signal():
    with pytest.raises(launcher.ProcessExitedException) as ei:
        launcher.spawn(H.child_killed, nprocs=2, grace_period=5)
    assert ei.value.signal_name == "SIGKILL"


def test_spawn_suspend_is_not_a_failure():
    pc = launcher.spawn(H.child_suspends, nprocs=2, grace_period=5)
    assert pc.suspended


def test_c_header_and_ctypes_signatures_cover_the_same_entry_points():
    """Every kernel entry point declared in include/pdb200.h has a ctypes signature (ops/_sigs.py) with the same
    number of parameters, and nothing is declared to ctypes that the header does not have."""
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "pdb200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    decl = {}
    for m in re.finditer(r"\b(?:int|size_t|const char\*)\s+(pdb_\w+)\s*\(([^;{]*?)\)\s*;", header, flags=re.S):
        args = m.group(2).strip()
        decl[m.group(1)] = 0 if args in ("", "void") else len([a for a in args.split(",") if a.strip()])
    sigs_src = open(os.path.join(root, "pytorch_distributed_b200", "ops", "_sigs.py")).read()
    sigs = {}
    for m in re.finditer(r'sig\(\s*"(pdb_\w+)"\s*,\s*\[(.*?)\]', sigs_src, flags=re.S):
        body = m.group(2).strip()
        sigs[m.group(1)] = 0 if not body else len([a for a in re.split(r",(?![^()]*\))", body) if a.strip()])
    for m in re.finditer(r'"(pdb_\w+)"\s*:\s*\[(.*?)\]', sigs_src, flags=re.S):        # the OPTIONAL (comm) table
        sigs[m.group(1)] = len([a for a in m.group(2).split(",") if a.strip()])
    assert len(decl) > 40 and len(sigs) > 40
    missing = sorted(set(sigs) - set(decl))
    assert not missing, f"ctypes signatures without a declaration in pdb200.h: {missing}"
    for name, n in sigs.items():
        assert decl[name] == n, f"{name}: header has {decl[name]} parameters, _sigs.py has {n}"
