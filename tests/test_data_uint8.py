"""Raw 8-bit loader output (normalised on the device by the native engine) must describe the same data as the
reference-style float output (image / 255, int64 mask): utils/dataloading.py:31-40 of the reference."""
import numpy as np
import torch

from distributedpytorch_b200.utils import data as D


def test_synthetic_raw_uint8_matches_float():
    a = D.SyntheticSegmentation(4, (96, 64), seed=3)
    b = D.SyntheticSegmentation(4, (96, 64), seed=3, raw_uint8=True)
    for i in range(4):
        fa, fb = a[i], b[i]
        assert fb["image"].dtype == torch.uint8 and fb["mask"].dtype == torch.uint8
        assert fa["image"].dtype == torch.float32 and fa["mask"].dtype == torch.int64
        assert torch.equal(fa["image"], fb["image"].float() / 255.0)
        assert torch.equal(fa["mask"], fb["mask"].long())


def test_file_dataset_raw_uint8_matches_float(tmp_path):
    img_dir, mask_dir = tmp_path / "imgs", tmp_path / "masks"
    D.write_synthetic_files(img_dir, mask_dir, 3, size=(96, 64), seed=1, image_ext=".png")
    a = D.build_dataset(img_dir, mask_dir, (48, 32))
    b = D.build_dataset(img_dir, mask_dir, (48, 32), raw_uint8=True)
    assert len(a) == len(b) == 3
    for i in range(3):
        fa, fb = a[i], b[i]
        assert fb["image"].dtype == torch.uint8 and fb["image"].shape == (3, 32, 48)
        np.testing.assert_allclose(fa["image"].numpy(), fb["image"].float().numpy() / 255.0, rtol=0, atol=1e-7)
        assert torch.equal(fa["mask"], fb["mask"].long())
