"""MNIST CNN with the ADAG trainer — the reference's ``ddl_mnist_aztk.py`` on this framework.

Mirrors /root/reference/ddl_mnist_aztk.py section by section (cluster set-up :44-75, data load :78-107, feature
pipeline :110-164, model :167-199, distributed evaluation :201-210, training + report :213-224).  Spark/AZTK and the
Azure blob are replaced by the in-repo launcher and an MNIST-shaped synthetic frame (there is no network here);
pass ``--train-csv/--test-csv`` to read the real ``mnist_train.csv`` / ``mnist_test.csv`` files instead.

    python examples/ddl_mnist.py --num-workers 8            # one rank per GPU
    python examples/ddl_mnist.py --num-workers 2 --cpu      # plumbing run on CPU (gloo)
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from distributeddeeplearningspark_b200.data import (DataFrame, DenseTransformer, MinMaxTransformer, OneHotTransformer,
                                                     ReshapeTransformer, VectorAssembler, mnist_like_frame)
from distributeddeeplearningspark_b200.models import Activation, Conv2D, Dense, Flatten, MaxPooling2D, Sequential
from distributeddeeplearningspark_b200.trainers import (ADAG, AccuracyEvaluator, AllreduceTrainer, ModelPredictor,
                                                         SingleTrainer)
from distributeddeeplearningspark_b200.data import LabelIndexTransformer


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--num-processes', type=int, default=2)       # reference: 2 processes per executor
    ap.add_argument('--num-executors', type=int, default=4)       # reference: 4 executors
    ap.add_argument('--num-workers', type=int, default=None)
    ap.add_argument('--trainer', default='adag', choices=['adag', 'allreduce', 'single'])
    ap.add_argument('--batch-size', type=int, default=16)
    ap.add_argument('--communication-window', type=int, default=5)
    ap.add_argument('--num-epoch', type=int, default=5)
    ap.add_argument('--train-rows', type=int, default=8192)
    ap.add_argument('--test-rows', type=int, default=2048)
    ap.add_argument('--train-csv', default=None)
    ap.add_argument('--test-csv', default=None)
    ap.add_argument('--cpu', action='store_true')
    a = ap.parse_args()
    if a.cpu:
        os.environ['DDLS_FORCE_CPU'] = '1'

    ## Setup the environment
    num_workers = a.num_workers or a.num_executors * a.num_processes
    print('Number of desired executors: ', a.num_executors)
    print('Number of desired processes per executor: ', a.num_processes)
    print('Total number of workers: ', num_workers)

    ## Load data
    if a.train_csv:
        raw_dataset_train = DataFrame.read_csv(a.train_csv)
        raw_dataset_test = DataFrame.read_csv(a.test_csv)
    else:
        raw_dataset_train = mnist_like_frame(a.train_rows, seed=0, as_columns=True)
        raw_dataset_test = mnist_like_frame(a.test_rows, seed=1, as_columns=True)

    ## Data transformation
    features = raw_dataset_train.columns
    features.remove('label')
    vector_assembler = VectorAssembler(inputCols=features, outputCol='features')
    dataset_train = vector_assembler.transform(raw_dataset_train)
    dataset_test = vector_assembler.transform(raw_dataset_test)

    nb_classes = 10
    encoder = OneHotTransformer(nb_classes, input_col='label', output_col='label_encoded')
    dataset_train = encoder.transform(dataset_train)
    dataset_test = encoder.transform(dataset_test)

    transformer = MinMaxTransformer(n_min=0.0, n_max=1.0, o_min=0.0, o_max=250.0,
                                    input_col='features', output_col='features_normalized')
    dataset_train = transformer.transform(dataset_train)
    dataset_test = transformer.transform(dataset_test)

    reshape_transformer = ReshapeTransformer('features_normalized', 'matrix', (28, 28, 1))
    dataset_train = reshape_transformer.transform(dataset_train)
    dataset_test = reshape_transformer.transform(dataset_test)

    dataset_train = dataset_train.select('features_normalized', 'matrix', 'label', 'label_encoded')
    dataset_test = dataset_test.select('features_normalized', 'matrix', 'label', 'label_encoded')

    dense_transformer = DenseTransformer(input_col='features_normalized', output_col='features_normalized_dense')
    dataset_train = dense_transformer.transform(dataset_train)
    dataset_test = dense_transformer.transform(dataset_test)

    training_set = dataset_train.repartition(num_workers)
    test_set = dataset_test.repartition(num_workers)
    training_set.cache()
    test_set.cache()

    ## Model and evaluation metric definition
    nb_filters, pool_size, kernel_size, input_shape = 32, (2, 2), (3, 3), (28, 28, 1)
    convnet = Sequential()
    convnet.add(Conv2D(nb_filters, (kernel_size[0], kernel_size[1]), padding='valid', input_shape=input_shape))
    convnet.add(Activation('relu'))
    convnet.add(Conv2D(nb_filters, (kernel_size[0], kernel_size[1])))
    convnet.add(Activation('relu'))
    convnet.add(MaxPooling2D(pool_size=pool_size))
    convnet.add(Flatten())
    convnet.add(Dense(225))
    convnet.add(Activation('relu'))
    convnet.add(Dense(nb_classes))
    convnet.add(Activation('softmax'))
    optimizer_convnet = 'adam'
    loss_convnet = 'categorical_crossentropy'
    convnet.summary()

    def evaluate_accuracy(model, test_set, features='matrix'):
        evaluator = AccuracyEvaluator(prediction_col='prediction_index', label_col='label')
        predictor = ModelPredictor(keras_model=model, features_col=features)
        transformer = LabelIndexTransformer(output_dim=nb_classes)
        test_set = test_set.select(features, 'label')
        test_set = predictor.predict(test_set)
        test_set = transformer.transform(test_set)
        return evaluator.evaluate(test_set)

    ## Model training and evaluation
    common = dict(keras_model=convnet, worker_optimizer=optimizer_convnet, loss=loss_convnet, batch_size=a.batch_size,
                  num_epoch=a.num_epoch, features_col='matrix', label_col='label_encoded')
    if a.trainer == 'adag':
        trainer = ADAG(num_workers=num_workers, communication_window=a.communication_window, **common)
    elif a.trainer == 'allreduce':
        trainer = AllreduceTrainer(num_workers=num_workers, **common)
    else:
        trainer = SingleTrainer(**common)
    trained_model = trainer.train(training_set)

    print('----------------------------------------------------')
    print('Training time: ', trainer.get_training_time())
    print('Accuracy: ', evaluate_accuracy(trained_model, test_set))
    print('Number of parameter server updates: ', trainer.parameter_server.num_updates)


if __name__ == '__main__':
    main()
