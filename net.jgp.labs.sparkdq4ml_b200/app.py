"""DataQuality4MachineLearningApp.start() (DataQuality4MachineLearningApp.java:37-155) on the B200 path.

Same steps in the same order as the lab: register the two UDFs under their names, load the CSV, rule 1 + filter,
rule 2 + filter, label alias, VectorAssembler, LinearRegression(maxIter 40, regParam 1, elasticNet 1).fit, transform,
training summary, predict(40).  The console `show()` calls of the lab are replaced by a returned dict.

`fused=True` runs the B200-native shape of the same pipeline: no filtered copy is written, the two rules are applied
inside the single aggregation pass of the fit (and inside the evaluation pass).
"""
from __future__ import annotations

import numpy as np

from . import dq as dqmod
from . import ml
from .runtime import Context


def parse_lab_csv(data: bytes):
    """The lab's CSV dialect (App:52-55: header=false, inferSchema=true): records split on \\r, \\n or \\r\\n,
    `_c0` integer, `_c1` double.  Host-side reference of the dialect; the app itself ingests on the GPU (Context.read_csv)."""
    text = data.decode("ascii").replace("\r\n", "\n").replace("\r", "\n")
    guest, price = [], []
    for rec in text.split("\n"):
        if not rec:
            continue
        a, b = rec.split(",")
        guest.append(int(a))
        price.append(float(b))
    return np.asarray(guest, dtype=np.int32), np.asarray(price, dtype=np.float64)


class DataQuality4MachineLearningApp:
    def __init__(self, ctx: Context | None = None, device: int = 0):
        self.ctx = ctx or Context(device)
        self.udf = dqmod.UDFRegistration()

    def start(self, csv_bytes: bytes, fused: bool = False, to_device: bool = True, feature_value: float = 40.0):
        ctx = self.ctx
        # DQ section (App:46-49)
        self.udf.register("minimumPriceRule", dqmod.MinimumPriceDataQualityUdf(), dqmod.DoubleType)
        self.udf.register("priceCorrelationRule", dqmod.PriceCorrelationDataQualityUdf(), dqmod.DoubleType)
        # load + rename (App:52-59)
        guest, price = ctx.read_csv(csv_bytes, to_device=to_device)  # GPU ingest (dq4ml_csv_parse)
        n_loaded = int(guest.shape[0])
        out = {"n_loaded": int(n_loaded)}
        if not fused:
            # 1) min price (App:68-78)
            price_no_min = dqmod.callUDF(ctx, self.udf, "minimumPriceRule", price)
            guest, price, n1 = dqmod.dq_filter(ctx, guest, price, dqmod.rule1_only_params())
            out["price_no_min"] = price_no_min
            out["n_after_rule1"] = int(guest.shape[0])
            # 2) correlated price (App:86-90)
            price_correct_correl = dqmod.callUDF(ctx, self.udf, "priceCorrelationRule", price, guest)
            guest, price, _ = dqmod.dq_filter(ctx, guest, price, dqmod.rule2_only_params())
            out["price_correct_correl"] = price_correct_correl
            out["n_after_rule2"] = int(guest.shape[0])
            df = ml.Frame(ctx, {"guest": guest, "price": price})
        else:
            df = ml.Frame(ctx, {"guest": guest, "price": price}).with_fused_dq("guest", "price")
        # ML section: label alias + assembler (App:101-113)
        df = df.withColumn("label", df.col("price"))
        df = ml.VectorAssembler().setInputCols(["guest"]).setOutputCol("features").transform(df)
        # App:120-126
        lr = ml.LinearRegression().setMaxIter(40).setRegParam(1).setElasticNetParam(1)
        model = lr.fit(df)
        # App:129-154
        transformed = model.transform(df)
        summary = model.summary()
        out.update({
            "survivor_guest": guest if not fused else None,
            "survivor_price": price if not fused else None,
            "n_fit_rows": model.num_rows,
            "prediction_column": transformed.col("prediction"),
            "numIterations": summary.totalIterations(),
            "objectiveHistory": summary.objectiveHistory(),
            "residuals": summary.residuals(),
            "RMSE": summary.rootMeanSquaredError(),
            "r2": summary.r2(),
            "intercept": model.intercept(),
            "coefficients": model.coefficients(),
            "regParam": model.getRegParam(),
            "tol": model.getTol(),
            "prediction": model.predict([feature_value]),
            "solver_used": model.solver_used,
            "model": model,
        })
        return out
