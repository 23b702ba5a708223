"""``src/lib/data_generator/base.py`` of the reference: an index-shuffling batch sequence.
(The reference subclasses ``tf.keras.utils.Sequence``; Keras is not part of this path, so this
is a plain Python sequence with the same members.)"""
import numpy as np


class DataGenerator:
    def __init__(self, dataset, batch_size=3, shuffle=True):
        self.dataset = dataset
        self.batch_size = batch_size
        self.shuffle = shuffle
        self.indices = self.get_valid_indices()
        self.on_epoch_end()

    def get_valid_indices(self):
        return np.arange(len(self.dataset))

    def on_epoch_end(self):
        if self.shuffle:
            np.random.shuffle(self.indices)

    def __len__(self):
        return len(self.indices) // self.batch_size

    def __getitem__(self, index):
        return self.get_data(self.indices[index * self.batch_size:(index + 1) * self.batch_size])

    def __iter__(self):
        for i in range(len(self)):
            yield self[i]

    def get_data(self, batch):
        raise NotImplementedError('This method should be implemented in a subclass.')
