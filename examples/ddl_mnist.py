"""MNIST CNN trained with the dist-keras-style trainers of this framework.

Same workload as the reference's MNIST script (/root/reference/ddl_mnist_aztk.py: cluster shape :44-75, data :78-107,
feature pipeline :110-164, network :167-199, distributed evaluation :201-210, training and report :213-224), written
as a small program of its own: a config object, one function per stage, and every distributed part (training,
prediction, accuracy) running through ``spark.run`` with one rank per GPU.  Spark/AZTK and the Azure blob are
replaced by the in-repo launcher and an MNIST-shaped synthetic frame (there is no network here); ``--train-csv`` /
``--test-csv`` read real ``mnist_train.csv`` / ``mnist_test.csv`` files instead.

    python examples/ddl_mnist.py --num-workers 8                 # ADAG, one rank per GPU
    python examples/ddl_mnist.py --trainer allreduce --num-workers 8
    python examples/ddl_mnist.py --num-workers 2 --cpu           # plumbing run on CPU (gloo)
"""
import argparse
import os
import sys
from dataclasses import dataclass
from typing import Optional

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from distributeddeeplearningspark_b200 import data as D
from distributeddeeplearningspark_b200 import models as M
from distributeddeeplearningspark_b200 import trainers as T


@dataclass
class Config:
    executors: int = 4                 # the reference cluster: 4 executors x 2 processes = 8 workers
    processes_per_executor: int = 2
    workers: Optional[int] = None      # overrides executors x processes
    trainer: str = 'adag'              # adag | allreduce | single
    batch_size: int = 16
    communication_window: int = 5
    epochs: int = 5
    classes: int = 10
    pixel_max: float = 250.0           # the reference scales by 250, not 255 (normalised values reach 1.02)
    train_rows: int = 8192
    test_rows: int = 2048
    train_csv: Optional[str] = None
    test_csv: Optional[str] = None

    @property
    def num_workers(self) -> int:
        return self.workers or self.executors * self.processes_per_executor


def load(cfg: Config):
    """Raw frames with a ``label`` column and 784 pixel columns."""
    if cfg.train_csv:
        return D.DataFrame.read_csv(cfg.train_csv), D.DataFrame.read_csv(cfg.test_csv)
    return (D.mnist_like_frame(cfg.train_rows, seed=0, as_columns=True),
            D.mnist_like_frame(cfg.test_rows, seed=1, as_columns=True))


def featurise(frame, cfg: Config):
    """pixels -> one 784-vector -> [0, 1] -> (28, 28, 1) image; label -> one-hot.  Partitioned for the workers."""
    pixel_cols = [c for c in frame.columns if c != 'label']
    stages = [
        D.VectorAssembler(inputCols=pixel_cols, outputCol='features'),
        D.OneHotTransformer(cfg.classes, input_col='label', output_col='label_encoded'),
        D.MinMaxTransformer(n_min=0.0, n_max=1.0, o_min=0.0, o_max=cfg.pixel_max, input_col='features',
                            output_col='features_normalized'),
        D.ReshapeTransformer('features_normalized', 'matrix', (28, 28, 1)),
    ]
    for stage in stages:
        frame = stage.transform(frame)
    frame = frame.select('matrix', 'label', 'label_encoded').repartition(cfg.num_workers)
    frame.cache()
    return frame


def build_network(cfg: Config):
    net = M.Sequential([
        M.Conv2D(32, (3, 3), padding='valid', input_shape=(28, 28, 1)), M.Activation('relu'),
        M.Conv2D(32, (3, 3)), M.Activation('relu'),
        M.MaxPooling2D(pool_size=(2, 2)),
        M.Flatten(),
        M.Dense(225), M.Activation('relu'),
        M.Dense(cfg.classes), M.Activation('softmax'),
    ])
    net.compile(optimizer='adam', loss='categorical_crossentropy')
    return net


def make_trainer(net, cfg: Config):
    shared = dict(keras_model=net, worker_optimizer=net.optimizer_name, loss=net.loss_name, batch_size=cfg.batch_size,
                  num_epoch=cfg.epochs, features_col='matrix', label_col='label_encoded')
    if cfg.trainer == 'adag':
        return T.ADAG(num_workers=cfg.num_workers, communication_window=cfg.communication_window, **shared)
    if cfg.trainer == 'allreduce':
        return T.AllreduceTrainer(num_workers=cfg.num_workers, **shared)
    return T.SingleTrainer(**shared)


def accuracy(net, frame, cfg: Config) -> float:
    """Predict -> arg-max -> compare, a partition per rank; only the (correct, total) counters cross ranks."""
    return T.evaluate_accuracy(net, frame, features='matrix', label='label', nb_classes=cfg.classes,
                               num_workers=cfg.num_workers if cfg.num_workers > 1 else None)


def parse() -> Config:
    ap = argparse.ArgumentParser(description=__doc__.split('\n')[0])
    ap.add_argument('--num-executors', type=int, default=Config.executors)
    ap.add_argument('--num-processes', type=int, default=Config.processes_per_executor)
    ap.add_argument('--num-workers', type=int, default=None)
    ap.add_argument('--trainer', default=Config.trainer, choices=['adag', 'allreduce', 'single'])
    ap.add_argument('--batch-size', type=int, default=Config.batch_size)
    ap.add_argument('--communication-window', type=int, default=Config.communication_window)
    ap.add_argument('--num-epoch', type=int, default=Config.epochs)
    ap.add_argument('--train-rows', type=int, default=Config.train_rows)
    ap.add_argument('--test-rows', type=int, default=Config.test_rows)
    ap.add_argument('--train-csv', default=None)
    ap.add_argument('--test-csv', default=None)
    ap.add_argument('--cpu', action='store_true', help='run the ranks on CPU with gloo (no GPU needed)')
    a = ap.parse_args()
    if a.cpu:
        os.environ['DDLS_FORCE_CPU'] = '1'
    return Config(executors=a.num_executors, processes_per_executor=a.num_processes, workers=a.num_workers,
                  trainer=a.trainer, batch_size=a.batch_size, communication_window=a.communication_window,
                  epochs=a.num_epoch, train_rows=a.train_rows, test_rows=a.test_rows, train_csv=a.train_csv,
                  test_csv=a.test_csv)


def main():
    cfg = parse()
    print(f'workers: {cfg.num_workers} ({cfg.executors} executors x {cfg.processes_per_executor} processes)')
    raw_train, raw_test = load(cfg)
    train, test = featurise(raw_train, cfg), featurise(raw_test, cfg)
    net = build_network(cfg)
    net.summary()
    trainer = make_trainer(net, cfg)
    trained = trainer.train(train)
    print('-' * 52)
    print('Training time: ', trainer.get_training_time(), ' (training loop on the slowest rank: ', trainer.get_loop_time(), ')')
    print('Accuracy: ', accuracy(trained, test, cfg))
    print('Number of parameter server updates: ', trainer.parameter_server.num_updates)
    for r, ph in enumerate(trainer.get_phase_times()[:1]):
        print(f'rank {r} host phases (s): ' + ', '.join(f"{k} {v['total_s']:.2f}" for k, v in ph.items()))


if __name__ == '__main__':
    main()
