"""SelfSupervisedSystem of main.py:28-434 with the label -> classifier loop fed from the device.

Phase 1 (build_dataset, main.py:63-190) is the hot path: for every image the 79 features, all strategies, their scores
and the first maximum come from batched engine calls on frames that were decoded ahead into pinned memory
(compat.build_dataset_rows + frame_io.py); dataset_building.csv, dataset.pkl and the <stem>_<best>.png files keep
the reference's schema and names (main.py:150-183).  Phase 2 (train_classifier, main.py:225-335) and predict
(main.py:398-434) keep scikit-learn exactly as the reference uses it — a StandardScaler and three small classifiers on
79-D vectors are not GPU work — and take their features from the batched device extractor.  Plots are the
reference's reporting layer and are not rebuilt (out of scope).
"""
from __future__ import annotations

import os
import pickle
from pathlib import Path

import numpy as np

from . import compat
from .features import extract_all_features_batch
from .frame_io import FrameReader


class SelfSupervisedSystem:
    def __init__(self, config=None, batch: int = 8, workers: int = 4):
        if config is None:
            from config import Config                      # the maintainer's config.py, as main.py:22 does
            config = Config()
        self.config = config
        self.dataset = []
        self.classifier = None
        self.scaler = None
        self.feature_names = []
        self.batch, self.workers = batch, workers
        self.config.create_folders()

    def collect_images(self):                              # main.py:44-61
        files = []
        for fmt in self.config.SUPPORTED_FORMATS:
            files.extend(Path(self.config.IMAGE_FOLDER).glob(f"*{fmt}"))
        print(f"\n找到 {len(files)} 張影像")
        return files

    def build_dataset(self):
        files = self.collect_images()
        if not files:
            return
        strategies = self.config.STRATEGIES
        print(f"定義了 {len(strategies)} 種增強策略")
        self.dataset, rows = compat.build_dataset_rows(files, strategies, self.config.QUALITY_WEIGHTS,
                                                       strategy_folder=self.config.STRATEGY_FOLDER,
                                                       batch=self.batch, workers=self.workers)
        if rows:
            import pandas as pd
            csv_path = os.path.join(self.config.REPORT_FOLDER, "dataset_building.csv")
            pd.DataFrame(rows).to_csv(csv_path, index=False)                       # main.py:186-189
            print(f"\n資料集記錄已儲存: {csv_path}")
        if self.dataset:
            path = os.path.join(self.config.MODEL_FOLDER, "dataset.pkl")
            with open(path, "wb") as f:
                pickle.dump(self.dataset, f)                                        # main.py:192-196
            print(f"資料集已儲存: {path}")
            print(f"成功處理 {len(self.dataset)} 張影像")
        else:
            print("錯誤: 沒有成功處理任何影像")

    def train_classifier(self):                            # main.py:225-335, scikit-learn as in the reference
        from sklearn.ensemble import GradientBoostingClassifier, RandomForestClassifier
        from sklearn.model_selection import cross_val_score, train_test_split
        from sklearn.preprocessing import StandardScaler
        from sklearn.svm import SVC
        if not self.dataset:
            print("錯誤: 資料集為空，請先執行 build_dataset()")
            return None
        X = np.array([item["features"] for item in self.dataset])
        y = np.array([item["best_strategy"] for item in self.dataset])
        unique_classes = np.unique(y)
        X_train, X_test, y_train, y_test = train_test_split(X, y, test_size=self.config.TEST_SIZE,
                                                            random_state=self.config.RANDOM_SEED, stratify=y)
        self.scaler = StandardScaler()
        Xs_train, Xs_test = self.scaler.fit_transform(X_train), self.scaler.transform(X_test)
        classifiers = {
            "RandomForest": RandomForestClassifier(**self.config.CLASSIFIERS["random_forest"]),
            "GradientBoosting": GradientBoostingClassifier(**self.config.CLASSIFIERS["gradient_boosting"]),
            "SVM": SVC(**self.config.CLASSIFIERS["svm"], probability=True),
        }
        results = {}
        for name, clf in classifiers.items():
            clf.fit(Xs_train, y_train)
            cv = cross_val_score(clf, Xs_train, y_train, cv=self.config.CV_FOLDS)
            results[name] = {"model": clf, "train_acc": clf.score(Xs_train, y_train), "test_acc": clf.score(Xs_test, y_test),
                             "cv_mean": cv.mean(), "cv_std": cv.std()}
            print(f"{name}: 訓練 {results[name]['train_acc']:.3f} 測試 {results[name]['test_acc']:.3f} CV {cv.mean():.3f}")
        best = max(results, key=lambda k: results[k]["test_acc"])
        self.classifier = results[best]["model"]
        with open(os.path.join(self.config.MODEL_FOLDER, "trained_model.pkl"), "wb") as f:
            pickle.dump({"classifier": self.classifier, "scaler": self.scaler, "best_model_name": best, "results": results,
                         "feature_dim": X.shape[1], "classes": unique_classes.tolist()}, f)
        return results

    def predict_batch(self, image_paths):
        """main.py:398-434 for many files: [(prediction, {class: probability})] in input order; unreadable files raise."""
        if self.classifier is None or self.scaler is None:
            raise ValueError("模型尚未訓練，請先執行 train_classifier()")
        paths = [Path(p) for p in image_paths]
        feats = {}
        reader = FrameReader(paths, batch=self.batch, workers=self.workers)
        try:
            for ps, frames in reader:
                f = extract_all_features_batch(frames).cpu().numpy()
                for p, row in zip(ps, f):
                    feats[p] = row
            if reader.failed:
                raise ValueError(f"無法讀取影像: {reader.failed[0]}")
        finally:
            reader.close()
        X = self.scaler.transform(np.stack([feats[p] for p in paths]))
        pred = self.classifier.predict(X)
        proba = self.classifier.predict_proba(X)
        classes = self.classifier.classes_
        return [(pred[i], dict(zip(classes, proba[i]))) for i in range(len(paths))]

    def predict(self, image_path):
        return self.predict_batch([image_path])[0]

    def run(self):                                         # main.py:436-457
        self.build_dataset()
        if not self.dataset:
            print("\n錯誤: 無法建立資料集")
            return
        self.train_classifier()
