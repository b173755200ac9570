from boofcv_b200.synth import blob_image, shifted_pair, descriptors, planted_descriptors  # noqa: F401
