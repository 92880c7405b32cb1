"""The python blocks of docs/guide/*.md are executed, block after block in one namespace per file, so the guides cannot
drift from the API."""
import glob
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GUIDES = sorted(glob.glob(os.path.join(ROOT, "docs", "guide", "*.md")))


@pytest.mark.parametrize("path", GUIDES, ids=[os.path.basename(p) for p in GUIDES])
def test_guide_snippets_run(path, tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    blocks = re.findall(r"```python\n(.*?)```", open(path).read(), flags=re.S)
    assert blocks, "guide without runnable blocks"
    ns = {"__name__": "__guide__"}
    for i, code in enumerate(blocks):
        try:
            exec(compile(code, f"{os.path.basename(path)}[block {i}]", "exec"), ns)
        except Exception as e:
            raise AssertionError(f"{os.path.basename(path)} block {i} failed: {type(e).__name__}: {e}\n{code}") from e
