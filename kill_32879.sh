ssh test@127.0.0.1 "kill -9 2412"
ssh test@127.0.0.1 "kill -9 2413"
kill -9 2109
