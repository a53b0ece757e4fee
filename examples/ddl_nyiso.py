"""NYISO load forecasting with GRU and LSTM — the reference's ``ddl_nyiso_aztk.py`` / ``ddl_nyiso_hdi.ipynb``.

Follows /root/reference/ddl_nyiso_aztk.py: hyper-parameters :79-93, load + schema :95-118, min-max :120-137,
lag/lead windows :139-150, assemble/reshape :152-177, split :179-198, GRU + ADAG(adagrad) :200-218, predict +
de-normalise + MAPE :220-246, LSTM + ADAG(adam) :248-274.  The NYISO CSV lives in an Azure blob that is not
reachable here, so an hourly synthetic series with the same columns stands in (``--csv`` reads a real file).

    python examples/ddl_nyiso.py --num-workers 4
    python examples/ddl_nyiso.py --num-workers 2 --cpu --epochs 2 --hours 1200
    python examples/ddl_nyiso.py --num-workers 4 --hours 11566 --cell lstm --report docs/ddl_nyiso_run.md

``--cell`` runs one of the notebook's two training cells (``gru`` = ddl_nyiso_hdi.ipynb cells up to :730, ``lstm`` = :772-934)
or ``all``; ``--report`` writes what the notebook keeps as recorded cell outputs — worker count, sample rows of every
pipeline stage, parameter counts, parameter-server updates, training seconds, MAPE — next to the values recorded in
the reference notebook, as a Markdown file.
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from distributeddeeplearningspark_b200.data import (DataFrame, MinMaxTransformer, ReshapeTransformer, VectorAssembler,
                                                     Window, lag, lead, nyiso_like_frame)
from distributeddeeplearningspark_b200.models import GRU, LSTM, Dense, Sequential
from distributeddeeplearningspark_b200.trainers import ADAG, DynSGD, ModelPredictor, get_MAPE


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--num-processes', type=int, default=2)
    ap.add_argument('--num-executors', type=int, default=2)
    ap.add_argument('--num-workers', type=int, default=None)
    ap.add_argument('--epochs', type=int, default=20)
    ap.add_argument('--hours', type=int, default=11688)
    ap.add_argument('--csv', default=None)
    ap.add_argument('--trainer', default='adag', choices=['adag', 'dynsgd'])
    ap.add_argument('--cell', default='all', choices=['all', 'gru', 'lstm'])
    ap.add_argument('--report', default=None, help='write a notebook-style run record (Markdown) here')
    ap.add_argument('--cpu', action='store_true')
    a = ap.parse_args()
    rec = []          # (cell title, recorded text) pairs of the run record
    if a.cpu:
        os.environ['DDLS_FORCE_CPU'] = '1'
    num_workers = a.num_workers or a.num_executors * a.num_processes
    print('Number of desired executors: ' + str(a.num_executors))
    print('Number of desired processes per executor: ' + str(a.num_processes))
    print('Total number of workers: ' + str(num_workers))
    rec.append(('Workers', f'Total number of workers: {num_workers}   (notebook: 4, ipynb:93)'))

    ## Define variables
    LEN_SEQ_IN, LEN_SEQ_OUT = 24, 1
    LEN_EXTRA_IN = LEN_SEQ_OUT
    LEN_TEST_DATA = 120
    N_UNITS = 128
    BATCH_SIZE, COM_WINDOW, N_EPOCHS = 32, 5, a.epochs

    ## Load data
    if a.csv:
        import pandas as pd
        pdf = pd.read_csv(a.csv)
        pdf['TimeStamp'] = pd.to_datetime(pdf['TimeStamp'].str[:19], format='%m/%d/%Y %H:%M:%S')
        raw_df = DataFrame.from_pandas(pdf)
    else:
        raw_df = nyiso_like_frame(a.hours, seed=0)
    df = raw_df.select(['TimeStamp', 'Name', 'HourAvgLoad', 'temperature'])
    df.show(3)
    rec.append(('Raw rows (ipynb:207-209)', df.to_string(3) + f'\nrows: {df.count()}'))

    # Normalize the data
    orig_min, orig_max = df.agg_min('temperature'), df.agg_max('temperature')
    df = MinMaxTransformer(n_min=0.0, n_max=1.0, o_min=orig_min, o_max=orig_max, input_col='temperature',
                           output_col='NormTemp', is_vector=False).transform(df)
    orig_min, orig_max = df.agg_min('HourAvgLoad'), df.agg_max('HourAvgLoad')
    df = MinMaxTransformer(n_min=0.0, n_max=1.0, o_min=orig_min, o_max=orig_max, input_col='HourAvgLoad',
                           output_col='NormLoad', is_vector=False).transform(df)
    df.show(3)
    rec.append(('Min-max normalised rows (ipynb:237-239)', df.to_string(3)))

    # Create input features and output targets
    wSpec = Window.partitionBy('Name').orderBy('TimeStamp')
    df = wSpec.sorted(df)
    for n_lag in range(LEN_SEQ_IN, 0, -1):
        df = lag(df, 'NormLoad', n_lag, wSpec, 'NormLoad_lag' + str(n_lag))
    for n_lag in range(1, LEN_SEQ_OUT + 1):
        df = lead(df, 'NormLoad', n_lag, wSpec, 'NormLoad_next' + str(n_lag))
        df = lead(df, 'HourAvgLoad', n_lag, wSpec, 'OrigNormLoad_next' + str(n_lag))
        df = lead(df, 'NormTemp', n_lag, wSpec, 'NormTemp_next' + str(n_lag))
    df = df.na.drop()
    df.show(1)

    features = ['NormLoad_lag' + str(n) for n in range(LEN_SEQ_IN, 0, -1)] + \
               ['NormTemp_next' + str(n) for n in range(1, LEN_SEQ_OUT + 1)]
    df = VectorAssembler(inputCols=features, outputCol='features').transform(df)
    df = ReshapeTransformer('features', 'feature_matrix', (LEN_SEQ_IN + LEN_EXTRA_IN, 1)).transform(df)
    df = VectorAssembler(inputCols=['NormLoad_next' + str(n) for n in range(1, LEN_SEQ_OUT + 1)],
                         outputCol='labels').transform(df)
    df = VectorAssembler(inputCols=['OrigNormLoad_next' + str(n) for n in range(1, LEN_SEQ_OUT + 1)],
                         outputCol='labels2').transform(df)
    df = ReshapeTransformer('labels', 'label_matrix', (LEN_SEQ_OUT, 1)).transform(df)
    df.show(1)

    # Partition the data into training and testing sets
    df_train = df.limit(df.count() - LEN_TEST_DATA)
    df_test = df.orderBy('TimeStamp', ascending=False).limit(LEN_TEST_DATA).orderBy('TimeStamp', ascending=True)
    cols = ['features', 'feature_matrix', 'labels', 'labels2', 'label_matrix']
    df_train = df_train.select(*cols).repartition(num_workers).cache()
    df_test = df_test.select(*cols).repartition(num_workers).cache()
    rec.append(('Train / test split (ipynb:478)', f'training windows: {df_train.count()}, test windows: {df_test.count()}, '
                f'partitions: {[len(df_train.partition_indices(i, num_workers)) for i in range(num_workers)]}'))

    inverse_transformer = MinMaxTransformer(n_min=orig_min, n_max=orig_max, o_min=0.0, o_max=1.0,
                                            input_col='prediction', output_col='prediction2', is_vector=True)
    Trainer = ADAG if a.trainer == 'adag' else DynSGD

    def fit_and_score(model, optimizer, tag):
        model.summary()
        trainer = Trainer(keras_model=model, worker_optimizer=optimizer, loss='mean_squared_error',
                          num_workers=num_workers, batch_size=BATCH_SIZE, communication_window=COM_WINDOW,
                          num_epoch=N_EPOCHS, features_col='feature_matrix', label_col='labels')
        trained = trainer.train(df_train)
        print('Number of parameter updates ' + str(trainer.parameter_server.num_updates))
        print('Total training time in seconds ' + str(trainer.get_training_time()))
        print('  of which the training loop on the slowest rank: ' + str(trainer.get_loop_time()))
        df_pred = ModelPredictor(keras_model=trained, features_col='feature_matrix').predict(df_test.limit(24))
        df_pred = inverse_transformer.transform(df_pred)
        df_pred.show(24)
        actual = [list(r) for r in np.asarray(df_pred['labels2'])]
        pred = [list(r) for r in np.asarray(df_pred['prediction2'])]
        mape = get_MAPE(actual, pred)
        print(f'{tag} MAPE on 24 test hours: {mape:.3f} %')
        ref = {'GRU': '1425 updates, 88.475 s, MAPE 2.809 % (ipynb:608-609, 730), 50,049 parameters (ipynb:549-553)',
               'LSTM': '1425 updates, 99.254 s, MAPE 3.587 % (ipynb:817-818, 934), 66,689 parameters (ipynb:772-776)'}[tag]
        rec.append((f'{tag} + ADAG({optimizer})',
                    f'parameters: {sum(p.numel() for p in model.parameters()):,}\n'
                    f'Number of parameter updates {trainer.parameter_server.num_updates}\n'
                    f'Total training time in seconds {trainer.get_training_time():.3f} (driver wall clock incl. process start; '
                    f'training loop on the slowest rank: {trainer.get_loop_time():.3f} s)\n'
                    f'MAPE on 24 test hours: {mape:.3f} %\n'
                    f'samples/s: {trainer.parameter_server.num_updates * COM_WINDOW * BATCH_SIZE / trainer.get_training_time():,.0f} '
                    f'by the wall clock, {trainer.parameter_server.num_updates * COM_WINDOW * BATCH_SIZE / max(trainer.get_loop_time(), 1e-9):,.0f} in the loop\n'
                    f'reference notebook (4 CPU workers, real NYISO data): {ref}'))
        return mape

    # Create the GRU network (ddl_nyiso_aztk.py:200-246)
    gru_model = Sequential()
    gru_model.add(GRU(N_UNITS, input_shape=(LEN_SEQ_IN + LEN_EXTRA_IN, 1)))
    gru_model.add(Dense(LEN_SEQ_OUT, activation='linear'))
    if a.cell in ('all', 'gru'):
        fit_and_score(gru_model, 'adagrad', 'GRU')

    # Create the LSTM network (ddl_nyiso_aztk.py:248-274)
    lstm_model = Sequential()
    lstm_model.add(LSTM(N_UNITS, input_shape=(LEN_SEQ_IN + LEN_EXTRA_IN, 1)))
    lstm_model.add(Dense(LEN_SEQ_OUT, activation='linear'))
    if a.cell in ('all', 'lstm'):
        fit_and_score(lstm_model, 'adam', 'LSTM')
    if a.report:
        import platform
        import torch
        where = torch.cuda.get_device_name(0) if (torch.cuda.is_available() and not a.cpu) else platform.processor() or 'cpu'
        with open(a.report, 'w') as f:
            f.write('# NYISO load forecast: run record\n\nThe counterpart of the recorded cell outputs of '
                    '`/root/reference/ddl_nyiso_hdi.ipynb`, produced by `examples/ddl_nyiso.py` '
                    f'(`{" ".join(sys.argv[1:])}`) on {num_workers} x {where}.  Synthetic hourly series '
                    '(the NYISO blob is not reachable), so MAPE is not comparable with the notebook; update counts are.\n')
            for title, text in rec:
                f.write(f'\n## {title}\n\n```\n{text}\n```\n')
        print('run record written to ' + a.report)


if __name__ == '__main__':
    main()
