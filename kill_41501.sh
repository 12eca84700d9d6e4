ssh test@127.0.0.1 "kill -9 9525"
ssh test@127.0.0.1 "kill -9 9526"
kill -9 8873
