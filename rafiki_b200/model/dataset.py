"""``dataset_utils`` of the plugin API -- the part the serving path uses.

``resize_as_images`` keeps the reference's contract (rafiki/model/dataset.py:68-78: list of N
2-D lists -> ``uint8 [N, S, S]``, Pillow resize with the pinned 5.3.0 default filter, NEAREST)
but does it as one array conversion and one index gather instead of a per-image PIL loop; the
same gather also exists on the GPU (``rfk_resize_decode_u8``).  The zip/CSV dataset loaders of
the reference (``:140-268``) are training-side I/O; ``load_dataset_of_image_files`` reads the
same zip format with PIL for the drop-in's ``train``/``evaluate``.
"""
import csv
import io
import os
import zipfile

import numpy as np


class InvalidDatasetProtocolException(Exception): pass
class InvalidDatasetTypeException(Exception): pass
class InvalidDatasetFormatException(Exception): pass


def nearest_index_table(n_in, n_out):
    """Pillow NEAREST source index per destination index: ``floor((dst + 0.5) * n_in / n_out)``,
    in exact integer arithmetic."""
    dst = np.arange(n_out, dtype=np.int64)
    return np.minimum(((2 * dst + 1) * n_in) // (2 * n_out), n_in - 1)


def queries_to_uint8(images):
    """Python lists (or arrays) of pixel values -> ``uint8`` array, wrapping modulo 256 as the
    reference's ``np.asarray(x, dtype=np.uint8)`` does under its pinned NumPy."""
    a = np.asarray(images)
    if a.dtype == np.uint8:
        return a
    if a.dtype.kind == 'f':
        a = a.astype(np.int64)
    return (a.astype(np.int64) & 0xFF).astype(np.uint8)


class ImageFilesDataset():
    """IMAGE_FILES dataset: a zip with ``images.csv`` (path,class) and the image files
    (format of rafiki/model/dataset.py:211-268)."""

    def __init__(self, dataset_path, image_size=None):
        self._images = []
        self._classes = []
        with zipfile.ZipFile(dataset_path) as z:
            with z.open('images.csv') as f:
                rows = list(csv.DictReader(io.TextIOWrapper(f, 'utf-8')))
            from PIL import Image
            for row in rows:
                with z.open(row['path']) as imf:
                    im = Image.open(io.BytesIO(imf.read()))
                    if image_size is not None:
                        im = im.resize(image_size, Image.NEAREST)
                    self._images.append(np.asarray(im))
                self._classes.append(int(row['class']))
        self.size = len(self._images)
        self.classes = len(set(self._classes))

    def __getitem__(self, index):
        return (self._images[index], self._classes[index])

    def __len__(self):
        return self.size


class ModelDatasetUtils():
    def load_dataset_of_image_files(self, dataset_uri, image_size=None):
        return ImageFilesDataset(self.download_dataset_from_uri(dataset_uri), image_size)

    def resize_as_images(self, images, image_size):
        """rafiki/model/dataset.py:68-78."""
        if len(images) == 0:
            return np.asarray([])
        a = queries_to_uint8(images)
        if image_size is None:
            return a
        rows = nearest_index_table(a.shape[1], int(image_size[1]))
        cols = nearest_index_table(a.shape[2], int(image_size[0]))
        if a.shape[1] == len(rows) and a.shape[2] == len(cols):
            return np.ascontiguousarray(a)
        return np.ascontiguousarray(a[:, rows][:, :, cols])

    def download_dataset_from_uri(self, dataset_uri):
        """Only local paths and ``file://`` URIs: there is no network on the serving box."""
        if dataset_uri.startswith('file://'):
            dataset_uri = dataset_uri[len('file://'):]
        if '://' in dataset_uri:
            raise InvalidDatasetProtocolException('only local dataset paths are supported: {}'.format(dataset_uri))
        if not os.path.exists(dataset_uri):
            raise InvalidDatasetFormatException('dataset not found: {}'.format(dataset_uri))
        return dataset_uri


dataset_utils = ModelDatasetUtils()
