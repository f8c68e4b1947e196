#!/usr/bin/env python
"""The flow of the reference's python-scripts/LSTM-TensorFlow-IO-Kafka/cardata-v1.py on the B200 path.

Same steps, same Keras calls — Kafka is replaced by a synthetic Confluent-framed Avro batch of the KSQL schema:

    kafka_dataset → substr(5) → decode_avro → normalize_fn → window(look_back) / skip(look_back) → zip → batch(1).take(1000)
    → fit(epochs=5) → predict on the next 200 windows with the OutputCallback (np.array2string per row)

Usage: python scripts/lstm_cardata_v1_b200.py [n_records]
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200_sensor as bs  # noqa: E402

keras = bs.keras


def main(n_records=1300, take=1000, predict_take=200, verbose=2):
    schema = bs.schemas.schema_b_json()
    rows = bs.synth.sensor_rows(n_records, seed=314, outlier_frac=0.0)
    data, offsets, _ = bs.synth.avro_b_batch(rows, seed=314)
    records = bs.decode.substr(bs.decode.RecordBatch(data, offsets), 5, -1)   # remove kafka framing (:25)
    feats, _, _, _ = bs.decode.decode_avro_matrix(records, schema)            # deserialize avro (:28-49)
    dataset = keras.normalize_fn(feats, *bs.schemas.norm_arrays())            # dataset.map(normalize_fn) (:157)

    features = 18
    look_back = 1
    batch_size = 1

    # create and fit the LSTM network (:164-176)
    model = keras.models.Sequential()
    model.add(keras.layers.LSTM(32, activation="relu", input_shape=(look_back, features), return_sequences=True))
    model.add(keras.layers.LSTM(16, activation="relu", return_sequences=False))
    model.add(keras.layers.RepeatVector(look_back))
    model.add(keras.layers.LSTM(16, activation="relu", return_sequences=True))
    model.add(keras.layers.LSTM(32, activation="relu", return_sequences=True))
    model.add(keras.layers.TimeDistributed(keras.layers.Dense(features)))
    model.compile(metrics=["accuracy"], loss="mean_squared_error", optimizer="adam")

    # dataset_x = window(look_back, shift=1), dataset_y = skip(look_back); zip(...).batch(batch_size).take(1000) (:184-191):
    # a view of the first take + look_back rows; its targets are the rows one step later
    train = keras.Windows(dataset[: take + look_back], look_back, shift=1)
    history = model.fit(train, epochs=5, batch_size=batch_size, shuffle=False, verbose=verbose).history
    print("Training complete")

    class OutputCallback(keras.callbacks.Callback):                 # :203-219 (KafkaOutputSequence → an in-memory dict)
        def __init__(self, batch_size):
            self._sequence = {}
            self._batch_size = batch_size

        def on_predict_batch_end(self, batch, logs=None):
            index = batch * self._batch_size
            for outputs in logs["outputs"]:
                for output in outputs:
                    message = np.array2string(output)
                    self._sequence[index] = message
                    index += 1

    # dataset_x.batch(batch_size).skip(1000).take(200) (:199)
    dataset_predict = keras.Windows(dataset[take: take + predict_take + look_back - 1], look_back, shift=1)
    output = OutputCallback(batch_size)
    predict = model.predict(dataset_predict, batch_size=batch_size, callbacks=[output])
    print("Predict complete")
    return history, predict, output._sequence


if __name__ == "__main__":
    h, p, seq = main(int(sys.argv[1]) if len(sys.argv) > 1 else 1300)
    print("loss per epoch:", [round(v, 5) for v in h["loss"]])
    print("predict", p.shape, "first message", seq[0][:60], "...")
