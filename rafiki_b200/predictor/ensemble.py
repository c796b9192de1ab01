"""``ensemble_predictions`` with the reference's contract (rafiki/predictor/ensemble.py:6-34),
the IMAGE_CLASSIFICATION mean computed by CUDA kernel 3 (``rfk_ensemble_mean``).

Edge cases kept: no trials or no queries -> ``[]``; any other task -> the first trial's
predictions unchanged; ndarray results become lists.  The kernel sums in float64 in trial order
and rounds once to float32, so each returned Python float is ``float32(reference value)``.
"""
import numpy as np

from ..constants import TaskType


def _simplify_predictions(predictions):
    if isinstance(predictions, np.ndarray):
        predictions = predictions.tolist()
    if isinstance(predictions, (list, tuple)):
        predictions = list(predictions)
        for i in range(len(predictions)):
            if isinstance(predictions[i], np.ndarray):
                predictions[i] = predictions[i].tolist()
    return predictions


def ensemble_predictions(predictions_list, task):
    if len(predictions_list) == 0 or len(predictions_list[0]) == 0:
        return []
    if task == TaskType.IMAGE_CLASSIFICATION:
        import torch
        from .. import engine
        engine._require_cuda()
        probs = [torch.as_tensor(np.ascontiguousarray(np.asarray(p, dtype=np.float32))).cuda()
                 for p in predictions_list]
        predictions = engine.ensemble_mean(probs).cpu().numpy()
    else:
        predictions = predictions_list[0]
    return _simplify_predictions(predictions)
