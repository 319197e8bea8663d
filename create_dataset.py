"""Dataset sampling script (reference surface: /root/reference/create_dataset.py:42-66).

``python create_dataset.py`` reads ``TRAIN_DIR/TRAIN_FILE`` (Herbarium ``metadata.json``),
samples ``N_IMAGES`` rows, splits 80/20, writes ``./data/{train,test}_sample.csv`` and
copies the images under ``./data/img/{train,test}/``.

``python create_dataset.py --synthetic [--regen N_TRAIN N_TEST]`` needs no Herbarium download: it renders the
stand-in JPEGs for the committed ``data/{train,test}_sample.csv`` (rows marked ``license == 999``) into
``TRAIN_DIR`` / ``TEST_DIR``, so ``MPB_SYNTHETIC=0 python main.py`` runs out of the box.
"""
import sys

import utils
from mpi_pytorch_b200.data.dataset import (create_dataframe, create_dataset, make_synthetic_index,  # noqa: F401
                                           read_json, render_index_images)

if __name__ == '__main__':
    if '--synthetic' in sys.argv:
        import pandas as pd
        if '--regen' in sys.argv:
            i = sys.argv.index('--regen')
            make_synthetic_index('./data', int(sys.argv[i + 1]), int(sys.argv[i + 2]), num_classes=100, hw=(100, 68),
                                 render=False)
        n = render_index_images(pd.read_csv(utils.TRAIN_CSV), utils.TRAIN_DIR)
        te = pd.read_csv(utils.TEST_CSV)
        n += render_index_images(te, utils.TRAIN_DIR) + render_index_images(te, utils.TEST_DIR)
        print("Rendered stand-in images for {} index rows".format(n))
    else:
        tr, te = create_dataset(utils.TRAIN_DIR, utils.TRAIN_FILE, utils.N_IMAGES)
        print("Created train & test samples: {} {}".format(tr, te))
