"""Shared fixtures. Every test that needs a GPU is marked @pytest.mark.gpu; everything else runs on CPU."""
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_manifest():
    return json.load(open(os.path.join(GOLDEN, "manifest.json")))


def load_kats():
    return json.load(open(os.path.join(GOLDEN, "kat.json")))


def corpus_bytes(entry) -> bytes:
    return open(os.path.join(GOLDEN, entry["file"]), "rb").read()


def corpus_ids():
    return [f"{e['group']}_{e['id']}" for e in load_manifest()["files"]]


def corpus_entries():
    return load_manifest()["files"]


@pytest.fixture(scope="session")
def manifest():
    return load_manifest()


def sample_dtype(container_bytes):
    return {1: np.int8, 2: np.int16, 4: np.int32}[container_bytes]
