"""Live HPO monitoring: ``ModelPlot``, ``ParamSpanWidget``, ``ModelController``, ``ModelTaskData``, ``ModelPlotTable``.

Same class names, constructor signatures and data flow as the reference ``hpo_widgets.py`` (plot :17-142, table +
polling :145-370, controller :373-407, containers :410-484), but the state machine is *headless*: every widget keeps
plain Python state and renders to text / HTML / SVG on demand (``render_text()``, ``_repr_html_()``), so it works in
a terminal, in tests and -- when ipywidgets is installed -- inside an ``Output`` widget.  No browser-side JavaScript
(bqplot / qgrid) and no graphics API is involved.

Fixed relative to the reference: Stop / Restart really abort and resubmit tasks (:352-364,386-391 were
stubs); history rows are appended once instead of once per poll (:270-279); list-valued parameters are stringified
for display (:184); the polling thread holds a lock and stops on ``close()`` (:230-233).
"""
import html
import threading
import time
import traceback

import numpy as np

COLORS = ["#1f77b4", "#ff7f0e", "#2ca02c", "#d62728", "#9467bd", "#8c564b", "#e377c2"]   # 7-colour cycle (:33)


class ModelPlotTable(object):
    """Column-major growable table: ``append_row(dict)``, ``append_column(name, vals)``, ``to_dict()``."""

    def __init__(self, column_names):
        self.column_names = list(column_names)
        self.columns = {c: [] for c in self.column_names}

    def num_rows(self):
        return len(self.columns[self.column_names[0]]) if self.column_names else 0

    def append_row(self, row):
        for c in self.column_names:
            self.columns[c].append(row.get(c))

    def append_column(self, name, vals):
        vals = list(vals)
        n = self.num_rows()
        if self.column_names and len(vals) != n:
            raise ValueError("column %r has %d values, table has %d rows" % (name, len(vals), n))
        if name not in self.columns:
            self.column_names.append(name)
        self.columns[name] = vals

    def to_dict(self):
        return {c: list(v) for c, v in self.columns.items()}


class ModelTaskData(object):
    """Per-trial container: a plot table (one row per epoch) + the latest status columns."""

    def __init__(self, plot_columns, status_columns):
        self.plot_table = ModelPlotTable(plot_columns)
        self.status = {c: None for c in status_columns}

    def num_data_rows(self):
        return self.plot_table.num_rows()

    def append_plot_data_row(self, row):
        self.plot_table.append_row(row)

    def get_plot_data(self):
        return self.plot_table.to_dict()


class ModelPlot(object):
    """Line + scatter plot of y-series against x, auto-rescaling.  ``update(data)`` takes a column dict."""

    def __init__(self, y, x=None, xlim=None, ylim=None, xlabel=None, ylabel=None, title=None):
        self.y = [y] if isinstance(y, str) else list(y)
        self.x = x
        self.xlim = list(xlim) if xlim else [0.0, 1.0]
        self.ylim = list(ylim) if ylim else [0.0, 1.0]
        self.xlabel, self.ylabel, self.title = xlabel or (x or "index"), ylabel or "", title or ""
        self.series = {name: ([], []) for name in self.y}

    def update(self, data):
        for name in self.y:
            ys = [v for v in data.get(name, []) if v is not None]
            xs = list(data[self.x])[:len(ys)] if self.x and self.x in data else list(range(len(ys)))
            self.series[name] = (xs, ys)
        self.resize_fig()

    def resize_fig(self):
        xs = [v for s in self.series.values() for v in s[0]]
        ys = [v for s in self.series.values() for v in s[1] if np.isfinite(v)]
        if xs:
            self.xlim = [min(self.xlim[0], min(xs)), max(self.xlim[1], max(xs))]
        if ys:
            self.ylim = [min(self.ylim[0], min(ys)), max(self.ylim[1], max(ys))]

    # ---- renderers
    def render_text(self, width=48):
        ticks = " .:-=+*#%@"
        lines = ["%s  [%s vs %s]" % (self.title, ",".join(self.y), self.xlabel)]
        lo, hi = self.ylim
        for name in self.y:
            ys = self.series[name][1]
            if not ys:
                lines.append("  %-10s (no data)" % name); continue
            spark = "".join(ticks[int(np.clip((v - lo) / (hi - lo + 1e-12), 0, 1) * (len(ticks) - 1))] for v in ys[-width:])
            lines.append("  %-10s %s  last=%.4f" % (name, spark, ys[-1]))
        return "\n".join(lines)

    def to_svg(self, w=480, h=260, pad=36):
        (x0, x1), (y0, y1) = self.xlim, self.ylim
        sx = lambda v: pad + (v - x0) / (x1 - x0 + 1e-12) * (w - 2 * pad)
        sy = lambda v: h - pad - (v - y0) / (y1 - y0 + 1e-12) * (h - 2 * pad)
        out = ['<svg xmlns="http://www.w3.org/2000/svg" width="%d" height="%d">' % (w, h),
               '<rect x="%d" y="%d" width="%d" height="%d" fill="none" stroke="#888"/>' % (pad, pad, w - 2 * pad, h - 2 * pad),
               '<text x="%d" y="16" font-size="12">%s</text>' % (pad, html.escape(self.title)),
               '<text x="%d" y="%d" font-size="11">%s</text>' % (w // 2, h - 6, html.escape(self.xlabel)),
               '<text x="4" y="%d" font-size="11">%s</text>' % (pad - 6, html.escape(self.ylabel))]
        for i, name in enumerate(self.y):
            xs, ys = self.series[name]
            col = COLORS[i % len(COLORS)]
            if ys:
                pts = " ".join("%.1f,%.1f" % (sx(a), sy(b)) for a, b in zip(xs, ys))
                out.append('<polyline points="%s" fill="none" stroke="%s" stroke-width="1.5"/>' % (pts, col))
                out += ['<circle cx="%.1f" cy="%.1f" r="2.5" fill="%s"><title>%s: %.4f</title></circle>' % (sx(a), sy(b), col, name, b) for a, b in zip(xs, ys)]
            out.append('<text x="%d" y="%d" font-size="11" fill="%s">%s</text>' % (w - pad - 70, pad + 14 * (i + 1), col, html.escape(name)))
        out.append("</svg>")
        return "".join(out)

    def _repr_html_(self):
        return self.to_svg()


class ModelController(object):
    """Owns the cluster client + load-balanced view; starts / stops / restarts trials and tracks their futures."""

    def __init__(self, ipp_cluster_id=None, client=None, retries=0):
        from ..cluster import Client
        self.client = client or Client(cluster_id=ipp_cluster_id)
        self.lview = self.client.load_balanced_view()
        self.lview.retries = retries
        self.futures = {}        # model_id -> AsyncResult
        self.specs = {}          # model_id -> (func, params)
        self.state = {}          # model_id -> 'running' | 'completed' | 'stopped' | 'failed'

    def start_model(self, model_id, func, params):
        self.specs[model_id] = (func, dict(params))
        self.futures[model_id] = self.lview.apply(func, **params)
        self.state[model_id] = "running"
        return self.futures[model_id]

    def stop_model(self, model_id):
        ar = self.futures.get(model_id)
        if ar is not None and self.state.get(model_id) == "running":
            ar.abort()
            self.state[model_id] = "stopped"

    def restart_model(self, model_id):
        self.stop_model(model_id)
        func, params = self.specs[model_id]
        return self.start_model(model_id, func, params)

    def set_model_completed(self, model_id, state="completed"):
        if self.state.get(model_id) == "running":
            self.state[model_id] = state

    def get_running_models(self):
        return {m: ar for m, ar in self.futures.items() if self.state.get(m) == "running"}

    def get_completed_models(self):
        return {m: ar for m, ar in self.futures.items() if self.state.get(m) == "completed"}

    def get_resource_usage(self):
        """Per-engine load (the reference left this as a stub, hpo_widgets.py:366-367)."""
        return self.client.queue_status()


class ParamSpanWidget(object):
    """One row per trial with columns ``status, epoch, <params...>, loss, val_loss, acc, val_acc``; a polling thread
    folds the dicts published by ``IPyParallelLogger`` into the table and the per-trial plots every ``interval`` s."""

    METRICS = ("loss", "val_loss", "acc", "val_acc")

    def _make_plot(self, vis_func, i):
        """One plot per trial, titled "Model i: {params}" as the reference does (hpo_widgets.py:218-220); a vis_func that takes no
        title still works."""
        title = "Model {}: {}".format(i, self.trial_params[i])
        try:
            return vis_func(title=title)
        except TypeError:
            return vis_func()

    def __init__(self, compute_func, vis_func, params, columns=None, ipp_cluster_id=None, output_layout=None,
                 qgrid_layout=None, client=None, interval=1.0, autostart=True):
        self.compute_func, self.vis_func = compute_func, vis_func
        self.param_names = list(params.keys())
        self.n = len(next(iter(params.values()))) if params else 0
        self.trial_params = [{k: self._py(params[k][i]) for k in self.param_names} for i in range(self.n)]
        self.columns = list(columns) if columns else ["status", "epoch"] + self.param_names + list(self.METRICS)
        self.rows = [dict({"status": "Not started", "epoch": None}, **{k: self._fmt(v) for k, v in tp.items()},
                          **{m: None for m in self.METRICS}) for tp in self.trial_params]
        self.model_plots = [self._make_plot(vis_func, i) for i in range(self.n)]
        self.model_data = [ModelTaskData(["epoch"] + list(self.METRICS), ["status", "epoch"]) for _ in range(self.n)]
        self.controller = ModelController(ipp_cluster_id, client=client)
        self.selected = []
        self.debug_log = []
        self._lock = threading.RLock()
        self._stop_updates = threading.Event()
        self.interval = interval
        self._thread = None
        if autostart:
            self.start_updates()

    @staticmethod
    def _py(v):
        if isinstance(v, np.ndarray):
            return v.tolist()
        if isinstance(v, np.generic):
            return v.item()
        return v

    @staticmethod
    def _fmt(v):
        return str(v) if isinstance(v, (list, tuple)) else v

    # ---- control
    def submit_computations(self):
        with self._lock:
            for i, tp in enumerate(self.trial_params):
                try:
                    self.controller.start_model(i, self.compute_func, tp)
                    self.rows[i]["status"] = "Submitted"
                except Exception:
                    self.debug_log.append(traceback.format_exc())

    def select(self, rows):
        with self._lock:
            self.selected = list(rows) if isinstance(rows, (list, tuple)) else [rows]

    def stop_selected_models(self, *_):
        with self._lock:
            for i in self.selected:
                self.controller.stop_model(i)
                self.rows[i]["status"] = "Stopped"

    def restart_selected_models(self, *_):
        with self._lock:
            for i in self.selected:
                self.model_data[i] = ModelTaskData(["epoch"] + list(self.METRICS), ["status", "epoch"])
                self.controller.restart_model(i)
                self.rows[i].update(status="Submitted", epoch=None, **{m: None for m in self.METRICS})

    def get_models_status(self):
        with self._lock:
            return [r["status"] for r in self.rows]

    # ---- polling
    def start_updates(self):
        if self._thread is None or not self._thread.is_alive():
            self._stop_updates.clear()
            self._thread = threading.Thread(target=self.update_data, kwargs={"interval": self.interval}, daemon=True, name="cib-psw-poll")
            self._thread.start()

    def update_data(self, interval=1.0):
        while not self._stop_updates.is_set():
            try:
                self.poll_once()
            except Exception:
                self.debug_log.append(traceback.format_exc())
            self._stop_updates.wait(interval)

    def poll_once(self):
        with self._lock:
            for i, ar in list(self.controller.get_running_models().items()):
                data = ar.data or {}
                hist = data.get("history") or {}
                md = self.model_data[i]
                epochs = hist.get("epoch", [])
                for j in range(md.num_data_rows(), len(epochs)):                 # append only NEW rows
                    md.append_plot_data_row(dict({"epoch": epochs[j]}, **{m: (hist.get(m) or [None] * len(epochs))[j] if j < len(hist.get(m) or []) else None
                                                                         for m in self.METRICS}))
                status = data.get("status")
                if status:
                    self.rows[i]["status"] = status
                if "epoch" in data:
                    self.rows[i]["epoch"] = data["epoch"]
                elif epochs:
                    self.rows[i]["epoch"] = epochs[-1]
                if epochs:
                    for m in self.METRICS:
                        if hist.get(m):
                            self.rows[i][m] = hist[m][-1]
                self.model_plots[i].update(md.get_plot_data())
                if ar.ready():
                    if ar.metadata["status"] == "ok":
                        self.rows[i]["status"] = "Ended Training"
                        self.controller.set_model_completed(i)
                    else:
                        self.rows[i]["status"] = "Failed: %s" % ((ar.metadata["error"] or ("?",))[0])
                        self.controller.set_model_completed(i, "failed")
                elif status == "Ended Training":
                    pass    # result message follows shortly

    def all_done(self):
        return not self.controller.get_running_models()

    def wait(self, timeout=None):
        t0 = time.time()
        while not self.all_done():
            if timeout is not None and time.time() - t0 > timeout:
                return False
            time.sleep(0.05)
        self.poll_once()
        return True

    def close(self):
        self._stop_updates.set()
        if self._thread is not None:
            self._thread.join(2.0)

    # ---- rendering
    def to_frame(self):
        import pandas as pd
        with self._lock:
            return pd.DataFrame([{c: r.get(c) for c in self.columns} for r in self.rows])

    def render_text(self):
        with self._lock:
            def cell(v):
                return "" if v is None else ("%.4f" % v if isinstance(v, float) else str(v))
            table = [self.columns] + [[cell(r.get(c)) for c in self.columns] for r in self.rows]
            widths = [max(len(str(row[j])) for row in table) for j in range(len(self.columns))]
            lines = ["  ".join(str(v).ljust(w) for v, w in zip(row, widths)) for row in table]
            lines.insert(1, "  ".join("-" * w for w in widths))
            for i in self.selected[:1]:
                lines += ["", self.model_plots[i].render_text()]
            return "\n".join(lines)

    def display_visualization(self, row=None):
        if row is not None:
            self.select(row)
        return self.model_plots[self.selected[0]] if self.selected else None

    def ipywidget(self, refresh=1.0):
        """Optional live view for a Jupyter front-end: an ``ipywidgets.HTML`` whose content is re-rendered every ``refresh``
        seconds from the headless state (table + plot of the selected trial).  ipywidgets is not a dependency: without it this
        raises ImportError and ``_repr_html_`` / ``render_text`` remain the way to look at the dashboard."""
        try:
            import ipywidgets
        except ImportError as e:
            raise ImportError("ParamSpanWidget.ipywidget() needs the optional 'ipywidgets' package; "
                              "use render_text() or _repr_html_() instead") from e
        view = ipywidgets.HTML(value=self._repr_html_())

        def loop():
            while not self._stop_updates.wait(refresh):
                view.value = self._repr_html_()
        threading.Thread(target=loop, daemon=True, name="cib-psw-view").start()
        return view

    def _repr_html_(self):
        with self._lock:
            head = "".join("<th>%s</th>" % html.escape(c) for c in self.columns)
            body = "".join("<tr>%s</tr>" % "".join("<td>%s</td>" % html.escape("" if r.get(c) is None else ("%.4f" % r[c] if isinstance(r[c], float) else str(r[c])))
                                                   for c in self.columns) for r in self.rows)
            plot = self.model_plots[self.selected[0]].to_svg() if self.selected else ""
            return "<table><thead><tr>%s</tr></thead><tbody>%s</tbody></table>%s" % (head, body, plot)
