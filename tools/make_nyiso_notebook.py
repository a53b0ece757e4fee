"""Build examples/ddl_nyiso.ipynb from examples/ddl_nyiso.py and a run record written by its --report option.

The reference ships its NYISO workload twice: as a script and as a notebook with recorded cell outputs
(/root/reference/ddl_nyiso_hdi.ipynb).  This tool cuts the example script at its stage comments into notebook cells and
attaches, as each cell's recorded output, the matching section of the run record of a REAL run of that script
(docs/ddl_nyiso_run.md: 4 x B200, 11,566 synthetic hours) — so the notebook's outputs are what this code printed.
    python tools/make_nyiso_notebook.py [docs/ddl_nyiso_run.md] [examples/ddl_nyiso.ipynb]"""
import json, os, re, sys, textwrap
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
record = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, 'docs', 'ddl_nyiso_run.md')
out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, 'examples', 'ddl_nyiso.ipynb')
src = open(os.path.join(ROOT, 'examples', 'ddl_nyiso.py')).read()

# ---- the run record: {section title: text inside its code fence}
sections = {}
for m in re.finditer(r'^## (.+?)\n\n```\n(.*?)\n```', open(record).read(), flags=re.S | re.M):
    sections[m.group(1).strip()] = m.group(2)
head = open(record).read().split('\n## ')[0]

# ---- the script body (inside main), dedented, without the argparse / report plumbing
body = src[src.index('    ## Define variables'):src.index("    if a.report:")]
body = textwrap.dedent(body)
body = '\n'.join(l for l in body.split('\n') if 'rec.append' not in l and not l.lstrip().startswith("f'") and not l.lstrip().startswith('f"'))
body = re.sub(r"\n\s+ref = \{.*?\}\[tag\]\n", "\n", body, flags=re.S)
cuts = [('## Define variables', None), ('## Load data', 'Raw rows (ipynb:207-209)'), ('# Normalize the data', 'Min-max normalised rows (ipynb:237-239)'),
        ('# Create input features and output targets', None), ('# Partition the data into training and testing sets', 'Train / test split (ipynb:478)'),
        ('inverse_transformer = ', None), ('# Create the GRU network', 'GRU + ADAG(adagrad)'), ('# Create the LSTM network', 'LSTM + ADAG(adam)')]
idx = [body.index(c[0]) for c in cuts] + [len(body)]
cells = []
md = lambda text: {'cell_type': 'markdown', 'metadata': {}, 'source': text.splitlines(keepends=True)}
def code(text, output, n):
    outs = [] if output is None else [{'name': 'stdout', 'output_type': 'stream', 'text': (output + '\n').splitlines(keepends=True)}]
    return {'cell_type': 'code', 'execution_count': n, 'metadata': {}, 'outputs': outs, 'source': text.strip('\n').splitlines(keepends=True)}

cells.append(md('# NYISO load forecasting with GRU and LSTM on B200 GPUs\n\nThe notebook form of `examples/ddl_nyiso.py` (the counterpart of the reference\'s '
                '`ddl_nyiso_hdi.ipynb`).  The code cells are the stages of that script; the recorded outputs are those of a real run of the script '
                '(`--num-workers 4 --hours 11566`, 4 x NVIDIA B200, synthetic hourly series: the NYISO blob is not reachable), taken from '
                '`docs/ddl_nyiso_run.md` by `tools/make_nyiso_notebook.py`.  MAPE is therefore not comparable with the reference notebook; '
                'the parameter-server update count (1425) is.\n'))
setup = '''import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.getcwd()))
from distributeddeeplearningspark_b200.data import (DataFrame, MinMaxTransformer, ReshapeTransformer, VectorAssembler,
                                                     Window, lag, lead, nyiso_like_frame)
from distributeddeeplearningspark_b200.models import GRU, LSTM, Dense, Sequential
from distributeddeeplearningspark_b200.trainers import ADAG, DynSGD, ModelPredictor, get_MAPE

class a:                       # what the script takes from its command line
    num_workers, epochs, hours, csv, trainer, cell = 4, 20, 11566, None, 'adag', 'all'
num_workers = a.num_workers
print('Total number of workers: ' + str(num_workers))'''
cells.append(code(setup, sections.get('Workers'), 1))
for k, (start, sec) in enumerate(cuts):
    cells.append(code(body[idx[k]:idx[k + 1]], sections.get(sec) if sec else None, k + 2))
nb = {'cells': cells, 'metadata': {'kernelspec': {'display_name': 'Python 3', 'language': 'python', 'name': 'python3'},
                                   'language_info': {'name': 'python'}}, 'nbformat': 4, 'nbformat_minor': 5}
json.dump(nb, open(out, 'w'), indent=1)
print(f'{out}: {len(cells)} cells, {sum(1 for c in cells if c.get("outputs"))} with recorded output')
