ssh test@127.0.0.1 "kill -9 3173"
ssh test@127.0.0.1 "kill -9 3174"
kill -9 2946
