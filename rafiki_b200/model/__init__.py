from .model import BaseModel, load_model_class, serve_model_class, \
    InvalidModelClassException, InvalidModelParamsException
from .knob import BaseKnob, CategoricalKnob, IntegerKnob, FloatKnob, FixedKnob, \
    serialize_knob_config, deserialize_knob_config
from .dataset import dataset_utils, ModelDatasetUtils, ImageFilesDataset
from .log import logger, ModelLogger
