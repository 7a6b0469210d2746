"""nkbs — B200-native nerfacto / semantic-NeRF-W ray-batch path behind the nerfstudio operator surface
(see DESIGN.md).  Importing the package does not need a GPU; calling an operator does."""
from .fields import (FieldHeadNames, HashEncoding, HashMLPDensityField, MLP, NerfactoField,  # noqa: F401
                     SceneContraction)
from .colormaps import apply_colormap, apply_depth_colormap, colormap_table  # noqa: F401
from .data import Cameras, RayGenerator, SemanticPalette, image_ray_indices, sample_masked_pixels  # noqa: F401
from .models import ModelConfig, NerfactoModel, SemanticNerfWModel  # noqa: F401
from .rays import Frustums, RayBundle, RaySamples  # noqa: F401
from .render import get_outputs_for_camera_ray_bundle, render_camera  # noqa: F401
from .renderers import (AccumulationRenderer, DepthRenderer, NearFarCollider, RGBRenderer,  # noqa: F401
                        SemanticRenderer, UncertaintyRenderer)
from .samplers import PDFSampler, ProposalNetworkSampler, UniformLinDispPiecewiseSampler  # noqa: F401

__version__ = "0.1.0"
