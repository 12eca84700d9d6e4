ssh test@127.0.0.1 "kill -9 5412"
kill -9 5044
