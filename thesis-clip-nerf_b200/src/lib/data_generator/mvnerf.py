"""``src/lib/data_generator/mvnerf.py`` of the reference: host-side assembly of the ray batch.

Pixel bookkeeping is integer work and must match the reference bit for bit: rays are (row, col)
pairs from ``bbox_biased_sample`` (numpy global RNG, 80 % inside the bbox which is the whole
image), ``u = col``, ``v = row``, targets are ``color[row, col, :3] / 255``."""
import numpy as np

from .base import DataGenerator
from .util import camera_parameters
from ..mvnerf.nerf_utils import bbox_biased_sample, get_specific_rays


class MVNeRFDataGenerator(DataGenerator):
    def __init__(self, dataset, n_rays_train=512, batch_size=1, n_views=2, **kwargs):
        super().__init__(dataset, batch_size, **kwargs)
        self.n_rays_train = n_rays_train
        self.n_views = n_views
        self.n_perspectives = self.dataset.datasets['color'].n_perspectives

    def generate_rays(self, color, camera_config):
        tgt_intrinsic = np.reshape(camera_config['intrinsics'], (3, 3)).astype(np.float32)
        h, w = color.shape[0], color.shape[1]
        rays = bbox_biased_sample(self.n_rays_train, np.array([0, 0, h, w]), h, w)
        r_o, r_d = get_specific_rays(rays[:, 1], rays[:, 0], camera_config['pose'], tgt_intrinsic)
        return r_d, r_o, rays

    @staticmethod
    def get_input(colors, camera_configs, r_d, r_o):
        e_invs, ks = zip(*[camera_parameters(c) for c in camera_configs])
        return (np.array([r_o], dtype=np.float32),
                np.array([r_d], dtype=np.float32),
                np.array([np.array(colors) / 255.0], dtype=np.float32),
                np.array([ks], dtype=np.float32),
                np.array([e_invs], dtype=np.float32))

    @staticmethod
    def get_target(color, rays):
        return np.array(color[rays[:, 0], rays[:, 1], :3]) / 255.0

    def get_data(self, batch):
        return self._assemble(batch, compact=False)

    def get_data_compact(self, batch):
        """``get_data`` with the source images left as the dataset's uint8 pixels ``[b, nv, H, W, C]`` (all stored
        channels): ``MVVNeRFRenderer.fit`` uploads those and evaluates the reference's ``colors / 255.0`` on the device
        (``cn_images_u8_to_f32``, bit-identical), which takes ~20 ms of numpy float conversion per 3 x 480 x 640 batch
        off the host.  Same numpy RNG draw sequence as ``get_data``."""
        return self._assemble(batch, compact=True)

    def compact_item(self, index):
        return self.get_data_compact(self.indices[index * self.batch_size:(index + 1) * self.batch_size])

    def _assemble(self, batch, compact):
        cols = [[] for _ in range(5)]
        targets = []
        color_ds, cam_ds = self.dataset.datasets['color'], self.dataset.datasets['camera_config']
        for i in batch:
            indices = np.random.choice(range(self.n_perspectives), size=self.n_views + 1, replace=False)
            src_indices, tgt_index = indices[:-1], indices[-1]
            tgt_color = color_ds.read_sample_at_idx(i, tgt_index)[..., :3]
            r_d, r_o, rays = self.generate_rays(tgt_color, cam_ds.read_sample_at_idx(i, tgt_index))
            targets.append(MVNeRFDataGenerator.get_target(tgt_color, rays))
            src_cfgs = [cam_ds.read_sample_at_idx(i, s) for s in src_indices]
            if compact:
                e_invs, ks = zip(*[camera_parameters(c) for c in src_cfgs])
                parts = ([r_o], [r_d], [np.stack([color_ds.read_sample_at_idx(i, s) for s in src_indices])], [ks], [e_invs])
            else:
                src_colors = [color_ds.read_sample_at_idx(i, s)[..., :3] for s in src_indices]
                parts = MVNeRFDataGenerator.get_input(src_colors, src_cfgs, r_d, r_o)
            for col, part in zip(cols, parts):
                col.extend(part)
        out = tuple(np.array(c, dtype=np.uint8 if (compact and k == 2) else np.float32) for k, c in enumerate(cols))
        return out, np.array(targets)
