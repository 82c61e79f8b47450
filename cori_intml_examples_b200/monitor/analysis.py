"""Post-hoc analysis used by the reference notebooks: accuracy / purity / efficiency with optional per-event weights,
ROC curve + AUC, output histograms, loss/accuracy curves as arrays (DistTrain_rpv.ipynb:342-362,378-396,447-514).
Pure NumPy (scikit-learn is optional); plotting is left to the user -- functions return arrays / dicts."""
import numpy as np


def summarize_metrics(labels, outputs, threshold=0.5, weights=None):
    """accuracy, purity (precision) and efficiency (recall), optionally weighted per event."""
    y = np.asarray(labels).ravel() > 0.5
    p = np.asarray(outputs).ravel() > threshold
    w = np.ones(y.shape, np.float64) if weights is None else np.asarray(weights, np.float64).ravel()
    tp = w[y & p].sum(); fp = w[~y & p].sum(); fn = w[y & ~p].sum(); tn = w[~y & ~p].sum()
    return {"accuracy": float((tp + tn) / max(w.sum(), 1e-300)), "purity": float(tp / max(tp + fp, 1e-300)),
            "efficiency": float(tp / max(tp + fn, 1e-300))}


def roc_curve(labels, outputs, weights=None):
    """(fpr, tpr, thresholds), thresholds descending -- same convention as sklearn.metrics.roc_curve."""
    y = (np.asarray(labels).ravel() > 0.5).astype(np.float64)
    s = np.asarray(outputs, np.float64).ravel()
    w = np.ones_like(s) if weights is None else np.asarray(weights, np.float64).ravel()
    order = np.argsort(-s, kind="mergesort")
    y, s, w = y[order], s[order], w[order]
    distinct = np.where(np.diff(s))[0]
    idx = np.r_[distinct, y.size - 1]
    tps = np.cumsum(y * w)[idx]; fps = np.cumsum((1 - y) * w)[idx]
    tps = np.r_[0, tps]; fps = np.r_[0, fps]
    return fps / max(fps[-1], 1e-300), tps / max(tps[-1], 1e-300), np.r_[s[idx[0]] + 1, s[idx]]


def auc(fpr, tpr):
    return float(getattr(np, "trapezoid", None)(tpr, fpr) if hasattr(np, "trapezoid") else np.trapz(tpr, fpr))


def output_histograms(labels, outputs, bins=20, weights=None):
    y = np.asarray(labels).ravel() > 0.5
    o = np.asarray(outputs).ravel()
    w = None if weights is None else np.asarray(weights).ravel()
    edges = np.linspace(0, 1, bins + 1)
    hs, _ = np.histogram(o[y], edges, weights=None if w is None else w[y])
    hb, _ = np.histogram(o[~y], edges, weights=None if w is None else w[~y])
    return {"edges": edges, "signal": hs, "background": hb}


def history_curves(history):
    """History (or its ``.history`` dict) -> {'epoch': [...], 'loss': [...], ...} as arrays."""
    h = getattr(history, "history", history)
    n = len(next(iter(h.values()))) if h else 0
    out = {k: np.asarray(v, float) for k, v in h.items()}
    out["epoch"] = np.asarray(getattr(history, "epoch", range(n)))
    return out
