"""Row b (drop-in boundary): the restated driver the GPU drop-in test runs (tests/helpers/train_lce_driver.py) makes
the same calls, with the same keyword names, in the same order as the reference's train/train_lce.py — checked
against the reference file itself whenever it is present (the build container; the GPU box has no /root/reference)."""
import ast
import inspect
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/train/train_lce.py"


def _calls(fn_node):
    """[(callee text, n positional args, sorted keyword names)] in source order, and the keys of dict literals."""
    out = []
    for node in ast.walk(fn_node):
        if isinstance(node, ast.Call):
            out.append((node.lineno, node.col_offset, ast.unparse(node.func), len(node.args),
                        tuple(sorted(k.arg for k in node.keywords if k.arg))))
        elif isinstance(node, ast.Dict):
            keys = tuple(k.value for k in node.keys if isinstance(k, ast.Constant))
            out.append((node.lineno, node.col_offset, "<dict>", 0, keys))
    return [c[2:] for c in sorted(out)]


def _main(path):
    tree = ast.parse(open(path).read())
    return next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "main")


@pytest.mark.skipif(not os.path.exists(REF), reason="reference tree not present on this box")
def test_driver_restates_reference_train_lce_call_for_call():
    ref, mine = _main(REF), _main(os.path.join(ROOT, "tests", "helpers", "train_lce_driver.py"))
    assert _calls(ref) == _calls(mine)
    ra, ma = ref.args, mine.args
    ref_params = [a.arg for a in ra.args]
    assert [a.arg for a in ma.args][:len(ref_params)] == ref_params       # + model_name (the local directory)
    assert [ast.unparse(d) for d in ra.defaults] == [ast.unparse(d) for d in ma.defaults][:len(ra.defaults)]


def test_lcemodel_signature_is_the_references():
    from pacednegatives_b200.lceT5 import LCEModel, LCEDataModule, LCEWeights, ChangeDifficulty
    sig = inspect.signature(LCEModel.__init__)
    assert list(sig.parameters)[:2] == ["self", "hparams"]
    # every other parameter is optional: LCEModel(hparams) alone is a valid call (lceT5.py:101-102)
    assert all(p.default is not inspect.Parameter.empty or p.kind == p.VAR_KEYWORD
               for n, p in list(sig.parameters.items())[2:])
    for name in ("pair_loss", "training_step", "configure_optimizers", "optimizers", "lr_schedulers",
                 "manual_backward", "log"):
        assert callable(getattr(LCEModel, name)), name
    assert list(inspect.signature(LCEDataModule.__init__).parameters)[1:] == [
        "data_dir", "corpus", "tokenizer", "batch_size", "shuffle", "use_max", "var", "n", "init_weight",
        "num_workers"]
    assert callable(ChangeDifficulty.on_train_batch_start) and callable(LCEWeights.forward)


def test_model_name_must_be_a_local_directory():
    """No hub access and no CPU fallback: a hub name fails loudly instead of training something else."""
    from pacednegatives_b200 import scorer
    with pytest.raises(FileNotFoundError):
        scorer.load_checkpoint("t5-base")
