ssh test@127.0.0.1 "kill -9 2604"
kill -9 2109
