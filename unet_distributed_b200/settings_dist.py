"""Constants of the reference's settings_dist.py (same names, same default values: settings_dist.py:1-42).
Paths can be overridden through the environment because the reference's are site specific."""
import os

BASE = os.environ.get("UNET_B200_BASE", "/home/bduser/unet/data/")
DATA_PATH = BASE + "/slices"
OUT_PATH = os.environ.get("UNET_B200_OUT_PATH", BASE + "slices/Results/")
IMG_ROWS = 128
IMG_COLS = 128
RESCALE_FACTOR = 1
SLICE_BY = 5

IN_CHANNEL_NO = 1
OUT_CHANNEL_NO = 1

EPOCHS = 10

BLOCKTIME = 0                # CPU threading knobs of the reference; unused on the GPU
NUM_INTRA_THREADS = 50
NUM_INTER_THREADS = 2
BATCH_SIZE = 1024            # GLOBAL batch (test_dist.py:390)

LEARNINGRATE = 0.0005
DECAY_STEPS = 100
LR_FRACTION = 0.2
CONST_LEARNINGRATE = True

USE_UPSAMPLING = False       # True = UpSampling2D decoder; False = Conv2DTranspose

MODEL_FN = "brainWholeTumor"
# MODE=1: FLAIR -> whole tumour; 2: T1 post -> active tumour; 3: T2 -> active core (settings_dist.py:30-33)
MODE = 1

# The parameter-server topology of the reference has no counterpart: replicas are the ranks torchrun starts.
PS_HOSTS = []
PS_PORTS = []
WORKER_HOSTS = ["127.0.0.1"]
WORKER_PORTS = ["29500"]

CHECKPOINT_DIRECTORY = os.environ.get("UNET_B200_CHECKPOINTS", "/tmp/unet_b200_checkpoints/")
TENSORBOARD_IMAGES = 3
