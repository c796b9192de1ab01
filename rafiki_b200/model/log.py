"""``logger`` object of the plugin API (interface of rafiki/model/log.py:14-192).

Training-time metric logging is outside the serving path; model files call it, so the methods
exist and forward to Python ``logging``.
"""
import logging

_log = logging.getLogger('rafiki_b200.model')


class ModelLogger():
    def define_loss_plot(self):
        self.define_plot('Loss Over Epochs', ['loss'], x_axis='epoch')

    def define_plot(self, title, metrics, x_axis=None):
        _log.debug('plot %s metrics=%s x_axis=%s', title, metrics, x_axis)

    def log_loss(self, loss, epoch):
        self.log(loss=loss, epoch=epoch)

    def log(self, msg='', **metrics):
        if msg:
            _log.info(msg)
        if metrics:
            _log.info('%s', metrics)


logger = ModelLogger()
