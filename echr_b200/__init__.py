"""echr_b200: the ECHR caption-generation hot path on B200 (sm_100a).  See DESIGN.md.

    CaptionGenerator                 the reference's module surface (caption_generator.py)
    engine.Engine / CaptionPipeline  batch-level calls over the C-ABI library (libechr_b200.so)
    sst_model.SST, proposals.*       the rows upstream of the path, output_stage.* downstream of it
"""
from .caption_generator import CaptionGenerator  # noqa: F401
