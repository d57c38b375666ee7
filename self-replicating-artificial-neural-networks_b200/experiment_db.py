"""SQLite results store with the reference's schema.

Same tables and columns as ``evolutionary_experiment/logic/experiment_db.py`` (``execution_info`` rewritten
on every start, ``generations`` and ``serann`` appended per generation, ``model_metrics``); column meanings are
documented in the reference ``README.md:57-137``.  The reference goes through sqlalchemy, which is not installed
here; pandas talks to a plain ``sqlite3`` connection just as well, so the files are interchangeable.
"""
from __future__ import annotations

import sqlite3
from contextlib import contextmanager

import numpy as np
import pandas as pd

INT_COLUMNS = ['genotype_size', 'initial_offspring_pool_size', 'max_serann_parameters', 'max_serann_tokens',
               'num_classification_classes', 'num_generations', 'num_seranns', 'offspring_pool_size_factor',
               'random_seed', 'training_batch_size', 'training_epochs']
FLOAT_COLUMNS = ['error_correction_probability', 'selection_pressure']


class ExperimentDB:
    def __init__(self, db_path, read_only: bool = False):
        """``read_only``: the ranks other than 0 of a multi-GPU run compute the same tables and must not write them."""
        self._db_path = str(db_path)
        self._read_only = read_only

    @property
    def db_path(self):
        return self._db_path

    @contextmanager
    def db_connection(self):
        conn = sqlite3.connect(self._db_path)
        try:
            yield conn
            conn.commit()
        except Exception as e:            # the reference prints and swallows DB failures (experiment_db.py:17-22)
            print(f'\033[91mFailed to access DB file\033[0m ({type(e).__name__}: {e})')
        finally:
            conn.close()

    def _table_exists(self, conn, name) -> bool:
        return conn.execute("select count(*) from sqlite_master where type='table' and name=?", (name,)).fetchone()[0] > 0

    def save_execution_info(self, start_time, parameters):
        if self._read_only:
            return
        with self.db_connection() as conn:
            row = {'start_time': str(start_time)}
            row.update(parameters)
            if 'ancestor_genotype' in parameters and parameters['ancestor_genotype'] is not None:
                row['ancestor_genotype'] = ''.join(str(i) for i in parameters['ancestor_genotype'])
            if 'classification_image_dimensions' in parameters:
                row['classification_image_dimensions'] = str(list(parameters['classification_image_dimensions']))
            existing = pd.read_sql('select * from execution_info', conn) if self._table_exists(conn, 'execution_info') else pd.DataFrame()
            existing = pd.concat([existing, pd.DataFrame([row])], ignore_index=True)
            existing.to_sql('execution_info', conn, index=False, if_exists='replace')

    def save_seranns_info(self, seranns_info: pd.DataFrame):
        if self._read_only:
            return
        with self.db_connection() as conn:
            genotype_str = seranns_info['genotype'].map(lambda x: str(np.asarray(x).tolist()))
            df = seranns_info.assign(genotype=genotype_str)
            for c in df.columns:              # sqlite has no boolean / object-bool type
                if df[c].dtype == bool:
                    df[c] = df[c].astype(int)
            df.to_sql('serann', conn, if_exists='append', index=True, index_label='id')

    def save_generation_info(self, generation_info: dict):
        if self._read_only:
            return
        with self.db_connection() as conn:
            row = {k: (str(v) if k == 'start_time' else v) for k, v in generation_info.items()}
            pd.DataFrame([row]).to_sql('generations', conn, index=False, if_exists='append')

    def save_model_metrics(self, model_name, metrics: pd.DataFrame):
        with self.db_connection() as conn:
            metrics.assign(model=model_name).to_sql('model_metrics', conn, index=False, if_exists='append')

    def get_last_execution_info(self):
        with self.db_connection() as conn:
            try:
                raw = pd.read_sql('select * from execution_info order by start_time desc limit 1', conn).iloc[0].copy()
            except Exception:
                return pd.DataFrame()
            for column in list(raw.keys()):
                if column in INT_COLUMNS and pd.notna(raw[column]):
                    raw[column] = int(float(raw[column]))
                elif column in FLOAT_COLUMNS and pd.notna(raw[column]):
                    raw[column] = float(raw[column])
            if 'classification_image_dimensions' not in raw and 'classification_image_height' in raw:
                raw['classification_image_dimensions'] = [int(raw['classification_image_height']), int(raw['classification_image_width'])]
            else:
                raw['classification_image_dimensions'] = [int(v) for v in eval(raw['classification_image_dimensions'])]
            if 'ancestor_genotype' in raw and isinstance(raw['ancestor_genotype'], str):
                raw['ancestor_genotype'] = [int(s) for s in raw['ancestor_genotype']]
            if 'initial_offspring_pool_size' not in raw:
                raw['initial_offspring_pool_size'] = 10
            if 'offspring_pool_size_factor' not in raw or raw['offspring_pool_size_factor'] == 1:
                raw['offspring_pool_size_factor'] = 3
            return raw
        return pd.DataFrame()

    def _count(self, table) -> int:
        with self.db_connection() as conn:
            if self._table_exists(conn, table):
                return int(conn.execute(f'select count(*) from {table}').fetchone()[0])
        return 0

    def get_generations_count(self) -> int:
        return self._count('generations')

    def get_executions_count(self) -> int:
        return self._count('execution_info')

    def get_serann_by_generation(self, generation: int) -> pd.DataFrame:
        with self.db_connection() as conn:
            df = pd.read_sql(f'select * from serann where generation = {int(generation)}', conn).set_index('id')
        df['genotype'] = df['genotype'].map(lambda x: np.array(eval(x.replace('nan', 'np.nan'))).astype('float16'))
        df['is_valid'] = df['is_valid'].astype(bool)
        return df
