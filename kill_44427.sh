ssh test@127.0.0.1 "kill -9 2619"
kill -9 2109
