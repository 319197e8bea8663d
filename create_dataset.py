"""Dataset sampling script (reference surface: /root/reference/create_dataset.py:42-66).

``python create_dataset.py`` reads ``TRAIN_DIR/TRAIN_FILE`` (Herbarium ``metadata.json``),
samples ``N_IMAGES`` rows, splits 80/20, writes ``./data/{train,test}_sample.csv`` and
copies the images under ``./data/img/{train,test}/``.
"""
import utils
from mpi_pytorch_b200.data.dataset import create_dataframe, create_dataset, read_json  # noqa: F401

if __name__ == '__main__':
    tr, te = create_dataset(utils.TRAIN_DIR, utils.TRAIN_FILE, utils.N_IMAGES)
    print("Created train & test samples: {} {}".format(tr, te))
