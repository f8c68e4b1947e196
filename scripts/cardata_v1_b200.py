#!/usr/bin/env python
"""The flow of the reference's python-scripts/AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.py on the B200 path.

Same steps, same Keras calls — Kafka is replaced by synthetic Confluent-framed Avro batches of the KSQL schema
(python-scripts/AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.avsc):

    kafka_dataset → substr(5) → decode_avro → normalize_fn → filter(label == "false") → fit → save → load_model
    → predict with the OutputCallback (np.array2string per row, keyed by a running index)

Usage: python scripts/cardata_v1_b200.py [n_records] [model.h5]
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200_sensor as bs  # noqa: E402

keras = bs.keras


def main(n_records=32 * 200, model_path="/tmp/path_to_my_model.h5", verbose=2):
    schema = bs.schemas.schema_b_json()
    # "kafka_dataset": one partition batch of Avro messages (cardata-v1.py:8-38)
    rows = bs.synth.sensor_rows(n_records, seed=314, outlier_frac=0.0)
    data, offsets, _ = bs.synth.avro_b_batch(rows, seed=314)
    records = bs.decode.RecordBatch(data, offsets)
    records = bs.decode.substr(records, 5, -1)                      # remove kafka framing (:13)
    feats, label, _, _ = bs.decode.decode_avro_matrix(records, schema)  # deserialize avro (:16-37) → f32[n,18], label codes

    # same autoencoder, 18 => 14 => 7 => 7 => 18 (:150-167)
    nb_epoch, batch_size = 5, 32
    input_dim, encoding_dim = 18, 14
    hidden_dim = int(encoding_dim / 2)
    learning_rate = 1e-7
    input_layer = keras.layers.Input(shape=(input_dim,))
    encoder = keras.layers.Dense(encoding_dim, activation="tanh", activity_regularizer=keras.regularizers.l1(learning_rate))(input_layer)
    encoder = keras.layers.Dense(hidden_dim, activation="relu")(encoder)
    decoder = keras.layers.Dense(hidden_dim, activation="tanh")(encoder)
    decoder = keras.layers.Dense(input_dim, activation="relu")(decoder)
    autoencoder = keras.models.Model(inputs=input_layer, outputs=decoder)
    autoencoder.compile(metrics=["accuracy"], loss="mean_squared_error", optimizer="adam")
    autoencoder.summary()
    autoencoder.set_normalization(*bs.schemas.norm_arrays())        # dataset.map(normalize_fn) (:179), fused into the kernels

    # only failure_occurred == "false" records train the model, 100 batches (:184-190)
    train = feats[label == bs.avro_schema.LABEL_FALSE][: batch_size * 100]
    history = autoencoder.fit(train, epochs=nb_epoch, batch_size=batch_size, shuffle=False, verbose=verbose).history
    print("Training complete")

    autoencoder.save(model_path)                                    # :199
    new_autoencoder = keras.models.load_model(model_path)           # :202
    new_autoencoder.set_normalization(*bs.schemas.norm_arrays())

    class OutputCallback(keras.callbacks.Callback):                 # :212-226 (KafkaOutputSequence → an in-memory dict)
        def __init__(self, batch_size):
            self._sequence = {}
            self._batch_size = batch_size

        def on_predict_batch_end(self, batch, logs=None):
            index = batch * self._batch_size
            for outputs in logs["outputs"]:
                for output in outputs:
                    self._sequence[index] = np.array2string(output)
                    index += 1

    output = OutputCallback(batch_size)
    predict_x = feats[batch_size * 100: batch_size * 200]           # .batch(32).skip(100).take(100) (:209-210)
    predict = new_autoencoder.predict(predict_x, batch_size=batch_size, callbacks=[output])
    print("Predict complete")
    mse, flag = new_autoencoder.score(predict_x, threshold=5)       # notebook scoring (nb2.0:1015,1163)
    print(f"records predicted: {predict.shape[0]}, messages produced: {len(output._sequence)}, flagged: {int(flag.sum())}, "
          f"final loss: {history['loss'][-1]:.4f}")
    return history, predict, output._sequence, mse, flag


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 32 * 200
    main(n, sys.argv[2] if len(sys.argv) > 2 else "/tmp/path_to_my_model.h5")
