"""Drop-in for the reference's use_trained_model.py (use_trained_model.py:13-295): EnhancementPredictor with the
reference's constructor and methods, batched on the B200; the network itself is the reference's own
ImprovedVGGParameterNet (vgg_16_UIE.py stays where it is, stock PyTorch)."""
import argparse
from pathlib import Path

import _bootstrap  # noqa: F401
from underwater_image_enhancement_b200.predictor import EnhancementPredictor  # noqa: F401

if __name__ == '__main__':
    parser = argparse.ArgumentParser(description='使用 vgg-16_UIE 模型進行影像增強')
    parser.add_argument('--input', type=str, required=True, help='輸入影像或資料夾')
    parser.add_argument('--output', type=str, required=True, help='輸出檔案或資料夾')
    parser.add_argument('--model', type=str, required=True, help='模型檔案（checkpoint）')
    parser.add_argument('--device', type=str, default='cuda', choices=['cuda', 'cpu'])
    args = parser.parse_args()
    predictor = EnhancementPredictor(model_path=args.model, device=args.device)
    input_path = Path(args.input)
    if input_path.is_file():
        predictor.process_single_image(args.input, args.output)
    elif input_path.is_dir():
        predictor.process_folder(args.input, args.output)
    else:
        raise ValueError("輸入路徑不存在或不是檔案/資料夾")
