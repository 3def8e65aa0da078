"""The reference's protobuf messages for a Python-3 runtime, and a ProxAffineProblem as the byte strings
`_solve.solve` takes.  TEST INFRASTRUCTURE (tests/ only).

The reference generates its Python classes with a 2016 protoc; those do not load in protobuf 6.  The messages are
built dynamically instead from the descriptor set `make -C oracle py3` lets the reference's own protoc write
(oracle/_ref/epsilon.desc; SURVEY.md section 8c names this route)."""
from __future__ import annotations

import os
import shutil
import tempfile
from typing import Dict, Optional, Tuple

HERE = os.path.dirname(os.path.abspath(__file__))
DESC = os.path.join(HERE, "_ref", "epsilon.desc")
MODULE_DIR = os.path.join(HERE, "_ref")

_classes = None


def available() -> bool:
    return os.path.isfile(DESC) and os.path.isfile(os.path.join(MODULE_DIR, "_solve.so"))


def messages():
    """{'Problem': class, 'Expression': class, 'SolverParams': class, 'SolverStatus': class, 'Constant': class}"""
    global _classes
    if _classes is None:
        from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
        fds = descriptor_pb2.FileDescriptorSet.FromString(open(DESC, "rb").read())
        pool = descriptor_pool.DescriptorPool()
        for fd in fds.file:
            pool.Add(fd)
        _classes = {name: message_factory.GetMessageClass(pool.FindMessageTypeByName(name))
                    for name in ("Problem", "Expression", "SolverParams", "SolverStatus", "Constant")}
    return _classes


def solve_inputs(problem, params: Optional[dict] = None) -> Tuple[bytes, bytes, Dict[str, bytes]]:
    """(Problem bytes, SolverParams bytes, {data_location: bytes}) of a ProxAffineProblem, through the same
    protobuf-text writer the compiled harness reads (oracle/ref_format.py)."""
    from google.protobuf import text_format
    from oracle.ref_format import write_reference_problem
    m = messages()
    tmp = tempfile.mkdtemp(prefix="epsproto_")
    try:
        write_reference_problem(problem, tmp, params)
        prob = text_format.Parse(open(os.path.join(tmp, "problem.pbtxt")).read(), m["Problem"]())
        par = text_format.Parse(open(os.path.join(tmp, "params.pbtxt")).read(), m["SolverParams"]())
        data = {}
        for name in os.listdir(os.path.join(tmp, "data")):
            data[name] = open(os.path.join(tmp, "data", name), "rb").read()
        return prob.SerializeToString(), par.SerializeToString(), data
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def prox_inputs(f, lam: float):
    """(Expression bytes, {data_location: bytes}) of one ProxFunction for `_solve.eval_prox`."""
    from google.protobuf import text_format
    from oracle.ref_format import write_reference_prox
    m = messages()
    tmp = tempfile.mkdtemp(prefix="epsproto_")
    try:
        write_reference_prox(f, lam, {}, tmp)
        expr = text_format.Parse(open(os.path.join(tmp, "f_expr.pbtxt")).read(), m["Expression"]())
        data = {}
        if os.path.isdir(os.path.join(tmp, "data")):
            for name in os.listdir(os.path.join(tmp, "data")):
                data[name] = open(os.path.join(tmp, "data", name), "rb").read()
        return expr.SerializeToString(), data
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def load_module():
    import importlib
    import sys
    if MODULE_DIR not in sys.path:
        sys.path.insert(0, MODULE_DIR)
    return importlib.import_module("_solve")
