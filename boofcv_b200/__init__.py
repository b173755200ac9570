"""boofcv_b200 — BoofCV's SURF detect / describe / associate path on B200 (sm_100a), behind the reference's own interfaces.

The package holds only what that path needs: csrc/ (hand-written CUDA kernels + the C ABI of include/boofsurf.h, built into
libboofsurf.so), the ctypes binding (_native), the stage-level calls (stages) and the mirror of the reference's factory /
interface layer (surf, associate).  There is no CPU fallback anywhere in this package."""
from .surf import (FactoryDetectDescribe, ConfigFastHessian, ConfigAverageIntegral, ConfigSlidingIntegral,  # noqa: F401
                   ConfigSurfDescribe, BrightFeature, AssociatedIndex, Point2D_F64, DetectDescribeSurf,
                   HostFeatureBuffers, ImageType, SurfPlanar_to_DetectDescribePoint)
from .associate import (FactoryAssociation, ScoreAssociateEuclideanSq_F64, MatchScoreType, AssociateGreedyGPU,  # noqa: F401
                        AssociateSurfBasic, WrapAssociateSurfBasic)
