"""Drop-in for the reference's ``src/lib/lmvnerf/model_v4.py``: ``LanguageNeRF`` is ``DeltaNGF`` with a
language-conditioned feature producer -- ``VisualFeatures`` + ``CombineCLIPVisualV4(use_dense=True, activation='elu')``
fed by the CLIP visual pyramid and the CLIP text embedding (reference :55-59, 178-187).

The two CLIP encoders are a frozen TensorFlow SavedModel outside the reference repository: they are plug-ins
(``clip_visual(images [(b n), H, W, 3]) -> 5-tuple``, ``clip_textual(tokens) -> [(b n), 1024]``).  With them given,
``encoders=True`` builds the torch producers (``src/lib/mvnerf/encoders.py``) and ``call`` runs the reference's chain;
otherwise ``model.feature_fn(images, clip_tokens) -> [(b n), H, W, 256]`` supplies the maps.  ``_call`` / ``infer`` /
``train_step`` are those of ``DeltaNGF`` (the sm_100a query path)."""
from ..delta_ngf.model import DeltaNGF, t_m_to_h_matrix, t_q_to_h_matrix  # noqa: F401
from ..mvnerf.model_v0 import _UpstreamStub


class LanguageNeRF(DeltaNGF):
    _groups = ('fine_embedding', 'visual_features', 'grasp_readout', 'combine_clip_visual')
    _backbone_groups = ('fine_embedding', 'visual_features', 'combine_clip_visual')

    def __init__(self, *args, encoders=False, clip_visual=None, clip_textual=None, lowres_features=False, **kwargs):
        device = kwargs.pop('device', None)
        super().__init__(*args, **kwargs)
        self.clip_visual = _UpstreamStub('CLIPVisualEncoder')
        self.clip_textual = _UpstreamStub('CLIPTextualEncoder')
        self.combine_clip_visual = _UpstreamStub('CombineCLIPVisualV4')
        for name, plug in (('clip_visual', clip_visual), ('clip_textual', clip_textual)):
            if plug is not None:
                del self._modules[name]
                setattr(self, name, plug)
        if encoders:
            from ..mvnerf.encoders import CombineCLIPVisualV4, VisualFeatures
            self.visual_features = VisualFeatures(256, self.original_image_size)
            self.combine_clip_visual = CombineCLIPVisualV4(use_dense=True, activation='elu',
                                                           final_upsample=not lowres_features)
        self._encoders = bool(encoders)
        self._clip_tokens = None
        if device is not None:
            self.to(device)

    def encode(self, src_images):
        b, n = src_images.shape[:2]
        flat = src_images.reshape(b * n, *src_images.shape[2:])
        if self.feature_fn is not None:
            feats = self.feature_fn(flat, self._clip_tokens)
        elif self._encoders:                                 # reference call(): :178-187
            text = self.clip_textual(self._clip_tokens)
            feats = self.combine_clip_visual((self.clip_visual(flat), self.visual_features(flat), text))
        else:
            return super().encode(src_images)
        return feats.reshape(b, n, *feats.shape[1:])

    def call(self, inputs, training=False, mask=None):
        self._clip_tokens = inputs[7] if len(inputs) > 7 else None
        return super().call(inputs, training=training)

    def train_step(self, data):
        self._clip_tokens = data[0][7] if len(data[0]) > 7 else None
        return super().train_step(data)
