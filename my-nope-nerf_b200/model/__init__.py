"""B200-native drop-in for the `model` package of js0n-lai/my-nope-nerf (hot path only).

`import model as mdl` resolves here when `my-nope-nerf_b200/` precedes the reference on sys.path; the exported
names match the reference's model/__init__.py:1-10.
"""
from model.checkpoints import CheckpointIO
from model.network import nope_nerf
from model.training import Trainer
from model.rendering import Renderer
from model.config import get_model
from model.official_nerf import OfficialStaticNerf
from model.poses import LearnPose
from model.intrinsics import LearnFocal
from model.eval_pose_one_epoch import Trainer_pose
from model.distortions import Learn_Distortion
from model.losses import Loss, Loss_Eval
from model.extracting_images import Extract_Images
from model.optim import FlatAdam
from model.scene import ResidentScene, BatchUploader
from model import parallel
