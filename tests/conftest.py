import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def golden_parser():
    return load_golden("parser_vectors.json")


@pytest.fixture(scope="session")
def golden_scorer():
    return load_golden("scorer_case.json")


@pytest.fixture(scope="session")
def golden_numbers():
    return load_golden("strtof_vectors.json")


@pytest.fixture(scope="session")
def criteo_lines():
    with open(os.path.join(GOLDEN, "criteo_excerpt.txt")) as f:
        lines = [l.rstrip("\n") for l in f]
    with open(os.path.join(GOLDEN, "criteo_excerpt_weights.txt")) as f:
        weights = [float(l) for l in f]
    return lines, weights
