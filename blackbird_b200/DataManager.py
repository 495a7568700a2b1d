"""DataManager mirror: the sqlite3 sink GenerateTrainingSamples writes to (DataManager.py:6-125).

Same database file (data/blackbird.db), same tables and columns, same method names, so games
written by this package are read back by the reference's trainer (Connection.GetGames ->
ExampleState.FromSerialized) and vice versa.  PutGames takes any number of serialized records in
ONE transaction; GenerateTrainingSamples hands it a whole batch of games at a time instead of one
commit per game.
"""
import gzip
import os
import sqlite3

_SCHEMA = """
CREATE TABLE ModelDim(ModelKey INTEGER PRIMARY KEY AUTOINCREMENT, GameType TEXT, Name TEXT, Version INTEGER DEFAULT 1);
CREATE TABLE ConfigurationDim(ConfigurationKey INTEGER PRIMARY KEY AUTOINCREMENT, ConfigJSON TEXT);
CREATE TABLE TrainingStatisticsFact(TrainingStatisticsKey INTEGER PRIMARY KEY AUTOINCREMENT, ModelKey INTEGER NOT NULL,
    OpponentKey INTEGER NOT NULL, Result INTEGER, Timestamp DATETIME DEFAULT CURRENT_TIMESTAMP,
    FOREIGN KEY (ModelKey) REFERENCES ModelDim(ModelKey), FOREIGN KEY (OpponentKey) REFERENCES ModelDim(ModelKey));
CREATE TABLE JobQueue(JobKey INTEGER PRIMARY KEY AUTOINCREMENT, QueueTime INTEGER, StartTime INTEGER, EndTime INTEGER,
    RPC INTEGER, PID INTEGER);
CREATE TABLE GameStateFact(GameStateKey INTEGER PRIMARY KEY AUTOINCREMENT, ModelKey INTEGER NOT NULL, GameType TEXT NOT NULL,
    State BYTES NOT NULL, FOREIGN KEY (ModelKey) REFERENCES ModelDim(ModelKey));
INSERT INTO ModelDim(Name, Version) VALUES('RANDOM', 0);
INSERT INTO ModelDim(Name, Version) VALUES('MCTS', 0);
"""
_MODEL_KEY = '(SELECT ModelKey FROM ModelDim WHERE Name = ? AND Version = ? ORDER BY ModelKey DESC LIMIT 1)'


class Connection(object):
    def __init__(self, isLocal=True, directory='data'):
        os.makedirs(directory, exist_ok=True)
        self._dir = directory
        self._conn = sqlite3.connect(os.path.join(directory, 'blackbird.db'))
        self.Cursor = self._conn.cursor()
        have = self.Cursor.execute("SELECT name FROM sqlite_master WHERE type='table' AND name='TrainingStatisticsFact';").fetchone()
        if have is None:                                  # DataManager.py:71-116
            self.Cursor.executescript(_SCHEMA)

    def GetLastVersion(self, gameType, name):
        """Newest Version of (name, gameType); registers version 1 when the model is unknown (DataManager.py:16-25)."""
        row = self.Cursor.execute('SELECT Version FROM ModelDim WHERE Name = ? AND GameType = ? ORDER BY Version DESC LIMIT 1;',
                                  (name, gameType)).fetchone()
        if row is None:
            self.PutModel(gameType, name, 1)
            return 1
        return row[0]

    def PutModel(self, gameType, name, version):
        self.Cursor.execute('INSERT INTO ModelDim(GameType, Name, Version) VALUES(?, ?, ?);', (gameType, name, version))
        self._conn.commit()

    def GetGames(self, name, version):
        rows = self.Cursor.execute('SELECT State FROM GameStateFact WHERE ModelKey = %s;' % _MODEL_KEY, (name, version))
        return [r[0] for r in rows.fetchall()]

    def PutGames(self, name, version, gameType, games):
        """games: iterable of serialized State records (DataManager.py:41-50)."""
        key = self.Cursor.execute('SELECT ModelKey FROM ModelDim WHERE Name = ? AND Version = ? ORDER BY ModelKey DESC LIMIT 1;',
                                  (name, version)).fetchone()
        key = key[0] if key else None                      # the reference's sub-select yields NULL -> IntegrityError
        self.Cursor.executemany('INSERT INTO GameStateFact(ModelKey, GameType, State) VALUES(?, ?, ?);',
                                ((key, gameType, g) for g in games))
        self._conn.commit()

    def DumpToZip(self, name, version):
        with gzip.open(os.path.join(self._dir, 'states_{0}_v{1}.gz'.format(name, version)), 'wb') as out:
            for state in self.GetGames(name, version):
                out.write(state + b'\x07\x07\x07')          # DataManager.py:52-55 framing

    def PutTrainingStatistic(self, result, name, version, opName, opVersion=0):
        one = '(SELECT ModelKey FROM ModelDim WHERE Name = ? AND Version = ?)'
        self.Cursor.execute('INSERT INTO TrainingStatisticsFact(ModelKey, OpponentKey, Result) VALUES(%s, %s, ?);' % (one, one),
                            (name, version, opName, opVersion, result))
        self._conn.commit()

    def Close(self):
        if self._conn is not None:
            self._conn.close()
            self._conn = None

    def __del__(self):
        try:
            self.Close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, exc_type, exc_val, exc_tb):
        self.Close()
