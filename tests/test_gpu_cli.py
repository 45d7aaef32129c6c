"""GPU: the CLI (`python -m data_free_quantization_b200.main_dfq`, flags of main_dfq.py:36-62) end to end -- DFQ steps,
then `inference_all` (main_dfq.py:66-113) through the eval plans: classification on the dynamic-range plan and on the
static u8 plan (--static), segmentation (--task seg, main_dfq.py:131-137) on the DeepLab plan, dynamic and static."""
import os

import pytest

pytestmark = pytest.mark.gpu
FLAGS = ["--quantize", "--relu", "--equalize", "--absorption", "--clip_weight", "--correction"]


@pytest.mark.parametrize("extra,images", [(["--batch", "8", "--batches", "2"], 16),
                                          (["--batch", "8", "--batches", "1", "--static"], 8),
                                          (["--task", "seg", "--batch", "1", "--batches", "1"], 1),
                                          (["--task", "seg", "--batch", "1", "--batches", "1", "--static"], 1)])
def test_cli_runs_the_pipeline_and_the_eval_plan(extra, images, capsys, tmp_path, monkeypatch):
    from data_free_quantization_b200 import main_dfq
    from data_free_quantization_b200.utils import quantize as Q
    monkeypatch.chdir(tmp_path)
    before = dict(Q.fused_calls)
    main_dfq.main(FLAGS + extra + ["--log"])
    out = capsys.readouterr().out
    assert "DFQ pipeline total" in out and f"({images} images" in out
    assert Q.fused_calls == before, "the CLI must evaluate through the plan, not through the per-module path"
    log = open(os.path.join(tmp_path, "dfq_result.txt")).read()
    assert "task: {}".format("seg" if "seg" in extra else "cls") in log and "Accuracy:" in log
