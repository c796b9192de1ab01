"""String enums of the reference that the serving path uses (rafiki/constants.py:59-62)."""


class TaskType():
    IMAGE_CLASSIFICATION = 'IMAGE_CLASSIFICATION'
    POS_TAGGING = 'POS_TAGGING'
    IMAGE_GENERATION = 'IMAGE_GENERATION'


class ModelDependency():
    # rafiki/constants.py (dependency names a model file may declare); kept so that
    # reference model files import unchanged
    TENSORFLOW = 'tensorflow'
    KERAS = 'Keras'
    SCIKIT_LEARN = 'scikit-learn'
    PYTORCH = 'torch'
