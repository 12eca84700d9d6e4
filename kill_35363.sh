ssh test@127.0.0.1 "kill -9 3473"
ssh test@127.0.0.1 "kill -9 3474"
kill -9 2946
