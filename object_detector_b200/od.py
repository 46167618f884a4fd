"""Mirror of the hot-path parts of pytoolkit/pytoolkit/dl/od.py: the anchor JSON
(``_VOC_JSON_DATA``, od.py:16-31), ``ObjectDetector.load_from_dict`` / ``load_voc`` (anchors only —
the CNN and its weights are out of scope), the ``encode_truth`` data-pipeline hook (od.py:169) and the
post-processing of ``predict`` (od.py:258-266)."""
import numpy as np
import torch

from . import ml, od_pb

_JSON_VERSION = '0.0.2'
_VOC_JSON_DATA = {
    "input_size": [320, 320],
    "map_sizes": [40, 20, 10],
    "num_classes": 20,
    "pb_size_patterns": [
        [1.2133152484893799, 1.6041910648345947],
        [2.301241874694824, 3.812246322631836],
        [3.2387468814849854, 7.3555989265441895],
        [5.467498779296875, 4.480413436889648],
        [4.758797645568848, 11.54628849029541],
        [10.89200496673584, 5.788760185241699],
        [7.736226558685303, 8.850168228149414],
        [4.47715950012207, 16.65940284729004]
    ],
    "version": "0.0.2"
}


class ObjectDetector:
    """Holds the prior boxes; the network itself stays in the caller's framework."""

    def __init__(self, input_size, map_sizes, num_classes):
        self.pb = od_pb.PriorBoxes(input_size, map_sizes, num_classes)
        self.strict_nms = True  # the reference builds the strict graph by default (od.py:80,207 -> od_net.py:175)

    def to_dict(self):
        data = {'version': _JSON_VERSION, 'input_size': self.pb.input_size, 'map_sizes': self.pb.map_sizes,
                'num_classes': self.pb.num_classes}
        data.update(self.pb.to_dict())
        return data

    @staticmethod
    def load_from_dict(data: dict):
        data = dict(data)
        if data['version'] == '0.0.1':  # od.py:70-71
            data.update(data.pop('pb'))
        od = ObjectDetector(input_size=data.get('input_size'), map_sizes=data.get('map_sizes'),
                            num_classes=data.get('num_classes'))
        od.pb.from_dict(data)
        return od

    @staticmethod
    def load_voc(input_size=(320, 320), strict_nms=True):
        """Anchor set of the published VOC models (od.py:80-97); weights are not loaded.  ``strict_nms``
        as in the reference signature (od.py:81): it selects nms_all_threshold=0.5 in the prediction tail
        (od_net.py:175) when ``predict_from_head`` is called without an explicit value."""
        input_size = tuple(input_size)
        assert input_size in ((320, 320), (640, 640))
        data = _VOC_JSON_DATA.copy()
        data['input_size'] = input_size
        od = ObjectDetector.load_from_dict(data)
        od.strict_nms = bool(strict_nms)
        return od

    def encode_truth(self, y):
        """The ProcessOutput hook of the training pipeline (od.py:169): one annotation -> (N, 3+C+4)."""
        return self.pb.encode_truth([y])[0]

    def predict_from_head(self, objs, clfs, locs, conf_threshold=0.01, top_k=200, nms_threshold=0.45, strict_nms=None):
        """What ``predict`` does after the CNN (od_net.py:186-190 + od.py:258-266, keep_aspect=False):
        head outputs -> list of ObjectsPrediction with confs >= conf_threshold.  ``strict_nms=True`` adds the
        all-class pass at 0.5 (od_net.py:175); the default ``None`` uses what ``load_voc`` was given, which itself
        defaults to True like the reference (od.py:80-97)."""
        if strict_nms is None:
            strict_nms = self.strict_nms
        out = self.pb.decode_nms(objs, clfs, locs, top_k=top_k, nms_threshold=nms_threshold,
                                 nms_all_threshold=0.5 if strict_nms else None)
        return unpack_predictions(out, conf_threshold)


def unpack_predictions(nms_out, conf_threshold=0.01):
    """od.py:258-266 for un-padded inputs (offset 0, size 1)."""
    p = nms_out.cpu().numpy() if isinstance(nms_out, torch.Tensor) else np.asarray(nms_out)
    res = []
    for row in p:
        mask = row[:, 1] >= conf_threshold
        res.append(ml.ObjectsPrediction(row[mask, 0].astype(np.int32), row[mask, 1], row[mask, 2:]))
    return res
