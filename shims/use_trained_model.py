"""Drop-in for the reference's use_trained_model.py (use_trained_model.py:13-295): EnhancementPredictor with the
reference's constructor and methods, batched on the B200; the network itself is the reference's own
ImprovedVGGParameterNet (vgg_16_UIE.py stays where it is, stock PyTorch)."""
import _bootstrap  # noqa: F401
from underwater_image_enhancement_b200.predictor import EnhancementPredictor  # noqa: F401


def _cli():
    """Same flags as use_trained_model.py:278-295 (--input file or folder, --output, --model, --device)."""
    import argparse
    import os
    p = argparse.ArgumentParser(description="vgg-16_UIE 模型影像增強 (B200)")
    for flag, text in (("--input", "輸入影像或資料夾"), ("--output", "輸出檔案或資料夾"), ("--model", "模型檔案（checkpoint）")):
        p.add_argument(flag, required=True, help=text)
    p.add_argument("--device", default="cuda", choices=["cuda", "cpu"])
    a = p.parse_args()
    if not os.path.exists(a.input):
        raise ValueError("輸入路徑不存在或不是檔案/資料夾")
    predictor = EnhancementPredictor(model_path=a.model, device=a.device)
    run = predictor.process_folder if os.path.isdir(a.input) else predictor.process_single_image
    run(a.input, a.output)


if __name__ == '__main__':
    _cli()
