"""``TfFeedForward`` -- drop-in for the reference model file
examples/models/image_classification/TfFeedForward.py, served by hand-written sm_100a kernels.

Same class name, knob configuration (``:20-28``) and ``BaseModel`` methods.  What changed is
underneath:

* ``predict`` (``:89-97``): the queries are uploaded as uint8 at the size they arrive in, resized
  on the GPU exactly as ``dataset_utils.resize_as_images`` does (Pillow NEAREST,
  ``rfk_resize_decode_u8``) and pushed through ``librafiki_b200`` (decode + BatchNormalization +
  Dense/ReLU layers + softmax on tcgen05 tensor cores with fp16 operands, or true fp32 with
  ``RAFIKI_B200_COMPUTE=f32``, or bf16 operands with ``=bf16``); the result is ``probs.tolist()`` as
  before.  Besides nested lists, ``predict`` takes a ``uint8`` ndarray ``[n, S, S(, ch)]`` (no
  per-pixel Python conversion); ``predict_array`` also returns an ndarray.
* ``dump_parameters`` / ``load_parameters`` (``:102-135``): the dict holds the Keras tensors as
  raw arrays (Dense kernels ``[in, out]``, biases, BatchNormalization gamma / beta / moving mean /
  moving variance) instead of a base64 HDF5 file, because neither TensorFlow nor h5py exists on
  the serving box.  A reference-made ``{'h5_model_base64': ...}`` dict is accepted through the
  built-in HDF5 reader (``rafiki_b200/keras_h5.py``).
* ``train`` / ``evaluate`` (``:38-87``) are not the accelerated path; ``train`` fits the same
  network with plain PyTorch autograd so that the plugin is complete, ``evaluate`` scores with
  the CUDA ``predict`` path.
"""
import os

import numpy as np

from rafiki_b200.model import BaseModel, InvalidModelParamsException, \
    IntegerKnob, CategoricalKnob, FloatKnob, FixedKnob, dataset_utils, logger
from rafiki_b200.constants import TaskType

PARAMS_FORMAT = 'rafiki_b200.TfFeedForward/1'
_ARRAY_KEYS = ('bn_gamma', 'bn_beta', 'bn_mean', 'bn_var')


class TfFeedForward(BaseModel):
    '''
    Implements a fully-connected feed-forward neural network with variable hidden layers
    for simple image classification
    '''
    @staticmethod
    def get_knob_config():
        return {
            'epochs': FixedKnob(2),
            'hidden_layer_count': IntegerKnob(1, 2),
            'hidden_layer_units': IntegerKnob(2, 128),
            'learning_rate': FloatKnob(1e-5, 1e-1, is_exp=True),
            'batch_size': CategoricalKnob([16, 32, 64, 128]),
            'image_size': FixedKnob(32)
        }

    def __init__(self, **knobs):
        super().__init__(**knobs)
        self._knobs = knobs
        self._params = None
        self._trial = None
        self._compute = os.environ.get('RAFIKI_B200_COMPUTE', 'f16')

    # ------------------------------------------------------------------ serving path
    def predict(self, queries):
        if len(queries) == 0:
            return []
        return self.predict_array(queries).tolist()

    def predict_array(self, queries):
        """Binary fast path: list or ``uint8`` ndarray ``[n, S, S(, ch)]`` -> ``float32`` ndarray ``[n, C]``."""
        if self._trial is None:
            raise InvalidModelParamsException('model has no parameters loaded')
        trial, x = self._queries_to_device(queries)
        return trial.forward(x).cpu().numpy()

    def predict_tensor(self, x):
        """Tensor fast path: ``uint8 [n, S, S(, ch)]`` CUDA tensor at the trial's ``image_size`` ->
        ``float32 [n, C]`` CUDA tensor."""
        if self._trial is None:
            raise InvalidModelParamsException('model has no parameters loaded')
        return self._trial.forward(x)

    def _queries_to_device(self, queries):
        """uint8 frames at their own size -> device.  Frames that are not the trial's ``image_size`` (the
        default knob is 32 for 28x28 queries; rafiki/model/dataset.py:76-78 resizes with Pillow NEAREST):
        tensor-core trials fold that resize into their first layer and read the frames as they are; the
        fp32 path resizes on the GPU (``rfk_resize_decode_u8``).  Returns (trial to run, device tensor)."""
        import torch
        from rafiki_b200 import engine
        from rafiki_b200.model.dataset import queries_to_uint8
        im_sz = self._image_size()
        frames = np.ascontiguousarray(queries_to_uint8(queries))
        x = torch.from_numpy(frames).cuda()
        trial = self._trial
        if frames.ndim >= 3 and (frames.shape[1] != im_sz or frames.shape[2] != im_sz):
            if frames.shape[1] != frames.shape[2]:
                raise ValueError('queries must be square images, got {}'.format(frames.shape[1:]))
            if self._compute in ('f32', 'fp32'):
                x = engine.resize_decode_u8(x, im_sz)
            else:
                trial = self._trial.for_input(frames.shape[1])
        return trial, x

    def _image_size(self):
        im_sz = self._knobs.get('image_size')
        if im_sz is None and self._params is not None:
            im_sz = self._params.get('image_size')
        return int(im_sz)

    def destroy(self):
        if self._trial is not None:
            self._trial.close()
            self._trial = None

    def dump_parameters(self):
        if self._params is None:
            raise InvalidModelParamsException('model is not trained')
        p = self._params
        params = {'format': PARAMS_FORMAT, 'image_size': int(p['image_size']), 'bn_eps': float(p.get('bn_eps', 1e-3))}
        for k in _ARRAY_KEYS:
            params[k] = np.asarray(p[k], np.float32)
        params['kernels'] = [np.asarray(k, np.float32) for k in p['kernels']]
        params['biases'] = [np.asarray(b, np.float32) for b in p['biases']]
        return params

    def load_parameters(self, params):
        if params is not None and params.get('h5_model_base64', None) is not None:
            from rafiki_b200 import keras_h5
            params = keras_h5.params_from_h5_base64(params['h5_model_base64'], self._knobs.get('image_size'))
        if params is None or any(params.get(k, None) is None for k in _ARRAY_KEYS + ('kernels', 'biases')):
            raise InvalidModelParamsException()
        from rafiki_b200.engine import Trial
        self._params = dict(params)
        self._params.setdefault('image_size', self._knobs.get('image_size'))
        self.destroy()
        self._trial = Trial(self._params, compute=self._compute)

    # ------------------------------------------------------------------ training side (plain PyTorch)
    def train(self, dataset_uri):
        import torch
        im_sz = self._knobs.get('image_size')
        bs = self._knobs.get('batch_size')
        ep = self._knobs.get('epochs')
        logger.define_loss_plot()
        dataset = dataset_utils.load_dataset_of_image_files(dataset_uri, image_size=[im_sz, im_sz])
        num_classes = dataset.classes
        images = np.asarray([dataset[i][0] for i in range(len(dataset))])
        classes = np.asarray([dataset[i][1] for i in range(len(dataset))])
        dev = torch.device('cuda' if torch.cuda.is_available() else 'cpu')
        net = self._build_model(int(np.prod(images.shape[1:])), num_classes).to(dev)
        opt = torch.optim.Adam(net.parameters(), lr=self._knobs.get('learning_rate'))
        X = torch.as_tensor(images.reshape(len(images), -1), dtype=torch.float32, device=dev)
        y = torch.as_tensor(classes, dtype=torch.long, device=dev)
        for epoch in range(ep):
            net.train()
            perm = torch.randperm(len(X), device=dev)
            total = 0.0
            for lo in range(0, len(X), bs):
                idx = perm[lo:lo + bs]
                if len(idx) < 2:
                    continue
                opt.zero_grad()
                loss = torch.nn.functional.cross_entropy(net(X[idx]), y[idx])
                loss.backward()
                opt.step()
                total += float(loss) * len(idx)
            logger.log_loss(total / max(len(X), 1), epoch)
        net.eval()
        bn = net[0]
        dense = [m for m in net if isinstance(m, torch.nn.Linear)]
        self._params = {
            'image_size': im_sz, 'bn_eps': 1e-3,
            'bn_gamma': bn.weight.detach().cpu().numpy(), 'bn_beta': bn.bias.detach().cpu().numpy(),
            'bn_mean': bn.running_mean.cpu().numpy(), 'bn_var': bn.running_var.cpu().numpy(),
            'kernels': [m.weight.detach().t().contiguous().cpu().numpy() for m in dense],
            'biases': [m.bias.detach().cpu().numpy() for m in dense],
        }
        self.load_parameters(self.dump_parameters())

    def evaluate(self, dataset_uri):
        im_sz = self._knobs.get('image_size')
        dataset = dataset_utils.load_dataset_of_image_files(dataset_uri, image_size=[im_sz, im_sz])
        images = [dataset[i][0] for i in range(len(dataset))]
        classes = np.asarray([dataset[i][1] for i in range(len(dataset))])
        probs = np.asarray(self.predict(images))
        return float(np.mean(np.argmax(probs, axis=1) == classes))

    def _build_model(self, d, num_classes):
        import torch
        units = self._knobs.get('hidden_layer_units')
        layers = self._knobs.get('hidden_layer_count')
        mods = [torch.nn.BatchNorm1d(d, eps=1e-3, momentum=0.01)]   # Keras: momentum 0.99, epsilon 1e-3
        width = d
        for _ in range(layers):
            mods += [torch.nn.Linear(width, units), torch.nn.ReLU()]
            width = units
        mods.append(torch.nn.Linear(width, num_classes))            # softmax lives in the loss / the kernel
        for m in mods:
            if isinstance(m, torch.nn.Linear):
                torch.nn.init.xavier_uniform_(m.weight)             # Keras glorot_uniform
                torch.nn.init.zeros_(m.bias)
        return torch.nn.Sequential(*mods)
