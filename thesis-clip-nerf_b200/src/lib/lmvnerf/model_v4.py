"""Drop-in for the reference's ``src/lib/lmvnerf/model_v4.py``: ``LanguageNeRF`` is ``DeltaNGF`` with a
language-conditioned feature producer (CLIP visual + textual encoders combined by ``CombineCLIPVisualV4``,
reference :55-59, 178-187).  The producer is upstream of the query path: set ``model.feature_fn`` to a callable
``(images [(b n), H, W, 3], clip_tokens) -> [(b n), H, W, 256]``; ``_call`` / ``infer`` are those of ``DeltaNGF``."""
from ..delta_ngf.model import DeltaNGF, t_m_to_h_matrix, t_q_to_h_matrix  # noqa: F401
from ..mvnerf.model_v0 import _UpstreamStub


class LanguageNeRF(DeltaNGF):
    _groups = ('fine_embedding', 'visual_features', 'grasp_readout', 'combine_clip_visual')
    _backbone_groups = ('fine_embedding', 'visual_features', 'combine_clip_visual')

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.clip_visual = _UpstreamStub('CLIPVisualEncoder')
        self.clip_textual = _UpstreamStub('CLIPTextualEncoder')
        self.combine_clip_visual = _UpstreamStub('CombineCLIPVisualV4')
        self._clip_tokens = None

    def encode(self, src_images):
        if self.feature_fn is None:
            return super().encode(src_images)
        b, n = src_images.shape[:2]
        feats = self.feature_fn(src_images.reshape(b * n, *src_images.shape[2:]), self._clip_tokens)
        return feats.reshape(b, n, *feats.shape[1:])

    def call(self, inputs, training=False, mask=None):
        self._clip_tokens = inputs[7] if len(inputs) > 7 else None
        return super().call(inputs, training=training)

    def train_step(self, data):
        self._clip_tokens = data[0][7] if len(data[0]) > 7 else None
        return super().train_step(data)
