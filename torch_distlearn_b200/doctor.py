"""``python -m torch_distlearn_b200.doctor`` — what this machine can run, in one screen.

Reports the interpreter / torch / CUDA versions, whether the two native libraries are built and load, and per GPU: name,
SM count, memory, the capabilities the symmetric heap needs (VMM, POSIX-fd handles, multicast = NVLS), and the
peer-access matrix.  ``--selftest`` additionally runs a one-rank collective on every backend that is available (the
CPU shared-memory transport always; the CUDA transport when a GPU is visible) and a tensor-core GEMM against float64.
Exit code 0 means every check that could run passed.  Nothing here is needed for training; it exists because the first
question after a failed launch is always "is it the build, the driver or the topology?".
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import time
import uuid
from typing import Any, Dict, List


def _lib_info() -> Dict[str, Any]:
    from . import _lib
    out: Dict[str, Any] = {}
    libdir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib")
    for key, name, loader in (("host", "libdistlearn_host.so", _lib.host), ("cuda", "libdistlearn_b200.so", _lib.cuda)):
        path = os.path.join(libdir, name)
        info: Dict[str, Any] = {"path": path, "built": os.path.exists(path)}
        if info["built"]:
            info["bytes"] = os.path.getsize(path)
            info["mtime"] = time.strftime("%Y-%m-%d %H:%M:%S", time.localtime(os.path.getmtime(path)))
            try:
                lib = loader()
                info["loads"] = True
                if key == "cuda":
                    info["version"] = int(lib.dl_version())
            except Exception as e:  # noqa: BLE001 - report, do not raise
                info["loads"] = False
                info["error"] = str(e)
        out[key] = info
    return out


def _gpu_info() -> List[Dict[str, Any]]:
    import torch
    if not torch.cuda.is_available():
        return []
    from . import _lib
    try:
        lib = _lib.cuda()
    except Exception:  # noqa: BLE001
        lib = None
    gpus = []
    n = torch.cuda.device_count()
    for d in range(n):
        p = torch.cuda.get_device_properties(d)
        g: Dict[str, Any] = {"index": d, "name": p.name, "sm": f"{p.major}.{p.minor}", "sms": p.multi_processor_count,
                             "memory_gb": round(p.total_memory / 2**30, 1)}
        if lib is not None:
            vmm, fdh, mc = C.c_int(0), C.c_int(0), C.c_int(0)
            if lib.dl_device_caps(d, C.byref(vmm), C.byref(fdh), C.byref(mc)) == 0:
                g.update(vmm=bool(vmm.value), fd_handles=bool(fdh.value), multicast=bool(mc.value))
        g["peers"] = [bool(torch.cuda.can_device_access_peer(d, q)) if q != d else True for q in range(n)]
        gpus.append(g)
    return gpus


def _selftest_host() -> Dict[str, Any]:
    import torch
    import torch_distlearn_b200 as dl
    comm = dl.HostCommunicator(0, 1, name=f"doctor-{os.getpid()}-{uuid.uuid4().hex[:6]}", heap_mb=4, timeout_s=20)
    try:
        model = torch.nn.Linear(16, 4)
        flat = dl.FlatParams(comm, model)
        sgd = dl.AllReduceSGD(comm, flat)
        before = flat.params.tensor.clone()
        model(torch.ones(2, 16)).sum().backward()
        want = before - 0.1 * flat.grads.tensor
        sgd.step(lr=0.1)
        ok = bool(torch.allclose(flat.params.tensor, want, atol=1e-6)) and float(flat.grads.tensor.abs().max()) == 0.0
        return {"ok": ok}
    finally:
        comm.close()


def _selftest_cuda() -> Dict[str, Any]:
    import torch
    import torch_distlearn_b200 as dl
    from .ops import tc_gemm
    dev = torch.device("cuda:0")
    comm = dl.P2PCommunicator(rank=0, world=1, device=dev, heap_mb=16)
    out: Dict[str, Any] = {}
    try:
        model = torch.nn.Linear(64, 8).to(dev)
        flat = dl.FlatParams(comm, model)
        sgd = dl.AllReduceSGD(comm, flat)
        before = flat.params.tensor.clone()
        model(torch.ones(4, 64, device=dev)).sum().backward()
        want = before - 0.1 * flat.grads.tensor
        sgd.step(lr=0.1)
        torch.cuda.synchronize()
        out["fused_sgd"] = bool(torch.allclose(flat.params.tensor, want, atol=1e-6))
        out["nvls"] = bool(getattr(comm, "has_multicast", False))
    finally:
        comm.close()
    if torch.cuda.get_device_capability(0)[0] >= 10:
        g = torch.Generator().manual_seed(0)
        a, b = torch.randn(256, 512, generator=g).to(dev), torch.randn(128, 512, generator=g).to(dev)
        got = tc_gemm(a, b)
        torch.cuda.synchronize()
        ref = a.double() @ b.double().T
        out["tc_gemm_rel_err"] = float((got.double() - ref).abs().max() / ref.abs().max())
        out["tc_gemm"] = out["tc_gemm_rel_err"] < 2e-5
    out["ok"] = all(v for k, v in out.items() if isinstance(v, bool) and k != "nvls")
    return out


def report(selftest: bool = False) -> Dict[str, Any]:
    import torch
    rep: Dict[str, Any] = {
        "python": sys.version.split()[0], "torch": torch.__version__, "torch_cuda": torch.version.cuda,
        "cuda_available": bool(torch.cuda.is_available()), "libraries": _lib_info(), "gpus": _gpu_info(),
    }
    rep["env"] = {k: v for k, v in os.environ.items() if k.startswith("DL_B200_")}
    problems: List[str] = []
    libs = rep["libraries"]
    if not libs["host"].get("built"):
        problems.append("CPU transport not built: python -m torch_distlearn_b200.build")
    elif not libs["host"].get("loads"):
        problems.append(f"CPU transport does not load: {libs['host'].get('error')}")
    if rep["cuda_available"]:
        if not libs["cuda"].get("built"):
            problems.append("CUDA library not built: python -m torch_distlearn_b200.build")
        elif not libs["cuda"].get("loads"):
            problems.append(f"CUDA library does not load: {libs['cuda'].get('error')}")
        for g in rep["gpus"]:
            if not g["sm"].startswith("10."):
                problems.append(f"GPU {g['index']} is sm_{g['sm']}: the kernels are compiled for sm_100a only")
            if not all(g["peers"]):
                problems.append(f"GPU {g['index']} cannot access every peer: the P2P transport needs a full NVLink/NVSwitch mesh")
            if g.get("vmm") is False:
                problems.append(f"GPU {g['index']} has no virtual-memory management: the heap falls back to cudaIpc (no NVLS)")
    if selftest:
        tests: Dict[str, Any] = {}
        if libs["host"].get("loads"):
            try:
                tests["host"] = _selftest_host()
            except Exception as e:  # noqa: BLE001
                tests["host"] = {"ok": False, "error": str(e)}
        if rep["cuda_available"] and libs["cuda"].get("loads"):
            try:
                tests["cuda"] = _selftest_cuda()
            except Exception as e:  # noqa: BLE001
                tests["cuda"] = {"ok": False, "error": str(e)}
        rep["selftest"] = tests
        for k, t in tests.items():
            if not t.get("ok"):
                problems.append(f"self-test '{k}' failed: {t}")
    rep["problems"] = problems
    return rep


def _print(rep: Dict[str, Any]) -> None:
    print(f"python {rep['python']}  torch {rep['torch']} (CUDA {rep['torch_cuda']})  cuda available: {rep['cuda_available']}")
    for key, info in rep["libraries"].items():
        state = "missing" if not info["built"] else ("loads" if info.get("loads") else f"DOES NOT LOAD ({info.get('error')})")
        extra = f"  {info.get('bytes', 0) // 1024} KiB, built {info.get('mtime')}" if info["built"] else ""
        print(f"  {key:4s} library: {state}{extra}")
    for g in rep["gpus"]:
        caps = ", ".join(f"{k}={'yes' if g.get(k) else 'no'}" for k in ("vmm", "fd_handles", "multicast") if k in g)
        peers = "".join("x" if p else "." for p in g["peers"])
        print(f"  GPU {g['index']}: {g['name']} sm_{g['sm']} {g['sms']} SMs {g['memory_gb']} GiB  {caps}  peers {peers}")
    if rep["env"]:
        print("  env: " + " ".join(f"{k}={v}" for k, v in sorted(rep["env"].items())))
    for k, t in rep.get("selftest", {}).items():
        print(f"  self-test {k}: {'ok' if t.get('ok') else 'FAILED'}  {json.dumps({a: b for a, b in t.items() if a != 'ok'})}")
    for p in rep["problems"]:
        print("  ! " + p)
    if not rep["problems"]:
        print("  no problems found")


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--selftest", action="store_true", help="run a one-rank fused step per backend and a tensor-core GEMM")
    ap.add_argument("--json", action="store_true", help="print the report as one JSON object")
    args = ap.parse_args(argv)
    rep = report(selftest=args.selftest)
    if args.json:
        print(json.dumps(rep))
    else:
        _print(rep)
    return 1 if rep["problems"] else 0


if __name__ == "__main__":
    sys.exit(main())
