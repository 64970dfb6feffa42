"""world_size-2 gloo check of the way bench.py combines ranks (replicas only):
barrier, max of the per-rank times, sum of the per-rank units.  CPU only."""
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import os, sys, json, time
    import torch, torch.distributed as dist
    sys.path.insert(0, %r)
    import bench
    rank, world, local = bench.dist_env()
    dist.init_process_group("gloo")
    dist.barrier()
    # each replica has its own seeded batch (seed = 1000 + rank), nothing is exchanged on the data path
    _, W, nl, reqs = bench.make_inputs(seed=1000 + rank, count=2048)
    my_ms = 10.0 * (rank + 1)
    t = torch.tensor([my_ms], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    units = torch.tensor([len(reqs)], dtype=torch.int64)
    dist.all_reduce(units, op=dist.ReduceOp.SUM)
    digest = torch.tensor([int(reqs["size"].astype("int64").sum())], dtype=torch.int64)
    gathered = [torch.zeros_like(digest) for _ in range(world)]
    dist.all_gather(gathered, digest)
    if rank == 0:
        print(json.dumps({"world": world, "max_ms": float(t.item()), "units": int(units.item()),
                          "digests": [int(g.item()) for g in gathered]}))
    dist.destroy_process_group()
""") % ROOT


def test_two_replicas_combine_with_max_time_and_summed_units(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29617")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                          "127.0.0.1", "--master-port", "29617", str(script)], capture_output=True, text=True, env=env,
                         timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    import json
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    r = json.loads(line)
    assert r["world"] == 2 and r["max_ms"] == 20.0 and r["units"] == 4096
    assert r["digests"][0] != r["digests"][1]  # different seeded batches per replica


def test_reference_arm_prints_one_json_line_on_cpu():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    import json
    r = json.loads(out.stdout.strip().splitlines()[-1])
    assert r["impl"] == "reference" and r["unit"] == "decisions/s" and r["value"] > 0
    assert r["cpu_baseline"]["kind"] == "port" and r["e2e"]["h2d_bytes_per_step"] == 0


def test_no_collective_after_the_other_ranks_have_left():
    """bench.py lets every rank but 0 return once the timed arms are done; a barrier or all_reduce after that point would
    wait for ranks that no longer exist (it did once: every N > 1 run hung until the NCCL watchdog)."""
    import ast
    src = open(os.path.join(ROOT, "bench.py")).read()
    tree = ast.parse(src)
    fn = [n for n in ast.walk(tree) if isinstance(n, ast.FunctionDef) and n.name == "run_ours"][0]
    leave = [n for n in fn.body if isinstance(n, ast.If) and "rank != 0" in ast.unparse(n.test)]
    assert len(leave) == 1
    after = [n for n in fn.body if n.lineno > leave[0].lineno]
    calls = [ast.unparse(c.func) for n in after for c in ast.walk(n) if isinstance(c, ast.Call)]
    bad = [c for c in calls if c in ("barrier", "dist.barrier", "dist.all_reduce", "dist.all_gather", "dist.broadcast")]
    assert not bad, bad
