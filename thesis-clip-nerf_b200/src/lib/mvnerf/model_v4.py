"""Drop-in for the reference's ``src/lib/mvnerf/model_v4.py``: ``MVVNeRFRenderer`` whose ``call`` produces
``combined_features`` itself -- ``VisualFeatures`` and ``CombineCLIPVisualV4`` as torch modules
(``encoders.py``; reference ``model_v4.py:29-31, 75-87``), the CLIP visual encoder as a plug-in (its SavedModel is
not part of the reference repository).  Everything after the feature maps is ``model_v0``'s sm_100a render path.

``lowres_features=True``: ``combine_clip_visual`` stops before its final 2x bilinear ``UpSampling2D`` and the gather
composes the upsample (``CnScene.feat_up = 2``) -- the same function of the same data with a 4x smaller map and
map gradient."""
import torch

from .encoders import CombineCLIPVisualV4, VisualFeatures
from .model_v0 import MVVNeRFRenderer as _RendererV0
from .model_v0 import render_view  # noqa: F401  (same free function)


class MVVNeRFRenderer(_RendererV0):
    def __init__(self, n_rays_train, n_rays_infer, n_views=2, n_samples=64, n_features=256, embed_direction_vector=True,
                 batch_size=1, near=0.7, far=1.5, original_image_size=(480, 640), lowres_features=False, clip_visual=None,
                 device=None, **kwargs):
        super().__init__(n_rays_train, n_rays_infer, n_views=n_views, n_samples=n_samples, n_features=n_features,
                         embed_direction_vector=embed_direction_vector, batch_size=batch_size, near=near, far=far,
                         original_image_size=original_image_size, device=None, **kwargs)
        if tuple(original_image_size) != (480, 640):
            raise ValueError('CombineCLIPVisualV4 hard-wires the 480 x 640 pyramid sizes (layers.py:597-616)')
        self.visual_features = VisualFeatures(n_features, original_image_size)
        self.combine_clip_visual = CombineCLIPVisualV4(final_upsample=not lowres_features)
        if clip_visual is not None:
            del self._modules['clip_visual']        # (the placeholder); any callable: images [(b n), H, W, 3] -> the CLIP 5-tuple
            self.clip_visual = clip_visual
        self.feature_fn = self._combined_features
        if device is not None:
            self.to(device)

    def encode(self, image):
        """reference ``encode`` (:45-48) is ``visual_features`` alone; the render path's feature plug-in is the whole chain"""
        return self._combined_features(self._to_dev(image))

    def _combined_features(self, images):
        clip_outputs = self.clip_visual(images)                        # (the reference's preprocess_tf lives in the plug-in)
        visual = self.visual_features(images)                          # [(b n), 240, 320, 256]
        text = torch.ones(images.shape[0], 1024, device=images.device)   # reference :82: clip_textuals = ones
        return self.combine_clip_visual((clip_outputs, visual, text))

    # ---- reference checkpoints (TensorFlow format) through tools/convert_tf_checkpoint.py
    def checkpoint_names(self):
        """the encoder variable names ``tools/convert_tf_checkpoint.py --names`` expects, one per line"""
        from .encoders import keras_variables
        return [n for g in ('visual_features', 'combine_clip_visual') for n in keras_variables(getattr(self, g), g + '.')]

    def load_reference_npz(self, path):
        """restore the encoders and the four NeRF sub-nets from the ``.npz`` the converter wrote"""
        import numpy as np
        from .encoders import load_keras_variables
        data = np.load(path)
        for g in ('visual_features', 'combine_clip_visual'):
            load_keras_variables(getattr(self, g), data, g + '.')
        with torch.no_grad():
            for net in ('coarse', 'fine'):
                for sub in ('embedding', 'readout'):
                    keys = sorted(k for k in data.files if k.startswith(f'{net}_{sub}/'))
                    flat = torch.cat([torch.as_tensor(data[k]).reshape(-1) for k in keys]).to(torch.float32)
                    mod = getattr(self, f'{net}_{sub}')
                    assert flat.numel() == mod.flat.numel(), (net, sub, flat.numel(), mod.flat.numel())
                    mod.flat.copy_(flat.to(mod.flat.device))
        return True
