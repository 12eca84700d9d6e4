ssh test@127.0.0.1 "kill -9 2745"
kill -9 2109
