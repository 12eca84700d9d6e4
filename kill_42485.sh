ssh test@127.0.0.1 "kill -9 3548"
ssh test@127.0.0.1 "kill -9 3549"
kill -9 2946
