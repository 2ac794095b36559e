"""bench.py, static checks that need no GPU.

A full multi-rank run (`torchrun ... bench.py --gpus N`, the driver's scaling run) executes the side measurements on rank 0
only while the other ranks wait at the last barrier.  Anything there that builds a process-group-aware object
(`perceiver.GraphedTrainStep` opens bucketed all-reduces when a process group is up) blocks forever: it hung rank 0 once.
Such calls must sit under `world == 1`."""
import ast
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SINGLE_RANK_ONLY = {"other_config_throughput", "reference_on_b200", "fp32_mode_step", "cpu_reference_arm"}


def _guards(test):
    src = ast.unparse(test)
    return "world == 1" in src or "not distributed" in src


def test_rank0_side_measurements_that_may_open_collectives_run_at_one_rank_only():
    tree = ast.parse(open(os.path.join(ROOT, "bench.py")).read())
    gpu_arm = next(n for n in ast.walk(tree) if isinstance(n, ast.FunctionDef) and n.name == "gpu_arm")
    found = set()

    def visit(node, guarded):
        if isinstance(node, ast.If):
            for child in node.body:
                visit(child, guarded or _guards(node.test))
            for child in node.orelse:
                visit(child, guarded)
            return
        if isinstance(node, ast.IfExp):
            visit(node.body, guarded or _guards(node.test))
            visit(node.orelse, guarded)
            visit(node.test, guarded)
            return
        if isinstance(node, ast.Call) and isinstance(node.func, ast.Name) and node.func.id in SINGLE_RANK_ONLY:
            found.add(node.func.id)
            assert guarded, f"bench.py:{node.lineno}: {node.func.id}() is reachable in a multi-rank run"
        for child in ast.iter_child_nodes(node):
            visit(child, guarded)

    visit(gpu_arm, False)
    assert found == SINGLE_RANK_ONLY, found


def test_default_invocation_is_one_gpu_and_short():
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert 'add_argument("--gpus", type=int, default=1)' in src
    assert 'add_argument("--steps", type=int, default=100)' in src


def test_result_line_carries_every_key_of_the_contract():
    tree = ast.parse(open(os.path.join(ROOT, "bench.py")).read())
    want = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
            "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline"}
    for fn, extra in (("gpu_arm", set()), ("reference_arm", {"impl"})):
        node = next(n for n in ast.walk(tree) if isinstance(n, ast.FunctionDef) and n.name == fn)
        keysets = [{k.value for k in d.keys if isinstance(k, ast.Constant)} for d in ast.walk(node) if isinstance(d, ast.Dict)]
        need = (want - {"clocks", "gpu_launches", "roofline"}) | extra if fn == "reference_arm" else want
        assert any(need <= ks for ks in keysets), (fn, sorted(need - max(keysets, key=lambda ks: len(need & ks))))
