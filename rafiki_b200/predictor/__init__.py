from .ensemble import ensemble_predictions
from .predictor import Predictor, PredictorTimeoutError
